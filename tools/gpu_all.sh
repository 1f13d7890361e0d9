#!/bin/bash
# kernel unit tests + model parity + per-op timing
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_kernels2.py -m gpu -q -x > gpurun_out/all_unit.log 2>&1; echo "unit rc=$?"; tail -3 gpurun_out/all_unit.log
timeout 900 python -m pytest tests/test_parity_r2.py tests/test_gpu_forward.py -m gpu -q -x > gpurun_out/all_parity.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/all_parity.log
timeout 300 python tools/time_ops.py > gpurun_out/all_timeops.log 2>&1; echo "time_ops rc=$?"; grep -E "us/launch|eager" gpurun_out/all_timeops.log
