#!/bin/bash
# ncu launch list (gpu__time_duration per launch, natural clocks) of the bench command, after a clean run of the same command;
# the weight-packing kernels of model set-up are filtered out by name
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-tiled > gpurun_out/ll_bench.json 2> gpurun_out/ll_bench.err; echo "bench rc=$?"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^(?!.*(pack_|to_f32_kernel|bias_l2)).*$' -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 1 --no-tiled > gpurun_out/ll_ncu.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/r02_launches.csv
