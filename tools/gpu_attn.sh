#!/bin/bash
# attention kernel iteration: unit parity, model parity, per-op timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "ocab_attention" > gpurun_out/attn_unit.log 2>&1; echo "unit rc=$?"; tail -3 gpurun_out/attn_unit.log
timeout 900 python -m pytest tests/test_parity_r2.py tests/test_gpu_forward.py -m gpu -q -x > gpurun_out/attn_parity.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/attn_parity.log
timeout 300 python tools/time_ops.py > gpurun_out/attn_timeops.log 2>&1; echo "time_ops rc=$?"; tail -30 gpurun_out/attn_timeops.log
