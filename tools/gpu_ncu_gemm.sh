#!/bin/bash
# ncu --set full of one fused-GEMM launch shape (aggr by default): HATB200_NO_TRACE plain run of tools/trace_gemm.py
mkdir -p gpurun_out
MODE=${1:-aggr}; CIN=${2:-180}; N=${3:-180}
HATB200_NO_TRACE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 2 -c 1 -f -o gpurun_out/r02_gemm_$MODE python tools/trace_gemm.py $CIN $N $MODE > gpurun_out/ncu_gemm_$MODE.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_gemm_$MODE.log
