#!/bin/bash
# ncu --set full of ONE launch of a kernel (name regex $1, skip $2) inside the bench command; CSV pages travel
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$1" -s ${2:-1} -c 1 -f -o /tmp/onek python bench.py --steps 2 --warmup 1 --no-tiled --no-cpu-baseline > gpurun_out/ncu_k.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/onek.ncu-rep --page raw --csv > gpurun_out/k_raw.csv 2>/dev/null
ncu -i /tmp/onek.ncu-rep --page source --csv > gpurun_out/k_src.csv 2>/dev/null
