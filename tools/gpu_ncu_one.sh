#!/bin/bash
# ncu --set full of ONE launch matching a kernel-name regex inside tools/time_ops.py; raw + source pages travel as CSV
mkdir -p gpurun_out
RX=${1:-FeatLast}; SKIP=${2:-1}
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$RX" -s $SKIP -c 1 -f -o /tmp/one python tools/time_ops.py conv_last > gpurun_out/ncu_one.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/one.ncu-rep --page raw --csv > gpurun_out/one_raw.csv 2>/dev/null
ncu -i /tmp/one.ncu-rep --page source --csv > gpurun_out/one_src.csv 2>/dev/null
ncu -i /tmp/one.ncu-rep --page source --print-source cuda,sass --csv > gpurun_out/one_src2.csv 2>/dev/null
ls -la gpurun_out/one_*.csv
