#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_tc -s 2 -c 1 -f -o gpurun_out/r02_attn python tools/time_ops.py ocab > gpurun_out/ncu_attn.log 2>&1; echo "ncu rc=$?"; tail -5 gpurun_out/ncu_attn.log
ls -la gpurun_out/*.ncu-rep
