#!/bin/bash
# fused-GEMM iteration: unit parity of every fused launch shape, model parity, per-op timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "linear or gemm or full_size" > gpurun_out/gemm_unit.log 2>&1; echo "unit rc=$?"; tail -3 gpurun_out/gemm_unit.log
timeout 900 python -m pytest tests/test_parity_r2.py tests/test_gpu_forward.py -m gpu -q -x > gpurun_out/gemm_parity.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/gemm_parity.log
HATB200_TRACE=1 timeout 300 python tools/time_ops.py > gpurun_out/gemm_timeops.log 2>&1; echo "time_ops rc=$?"; grep -E "gemm_tc\]" gpurun_out/gemm_timeops.log | sort | uniq -c; grep -E "us/launch|eager" gpurun_out/gemm_timeops.log
