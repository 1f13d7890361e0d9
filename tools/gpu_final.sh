#!/bin/bash
# round-end evidence on one GPU: full GPU test suite, smoke(), bench (both arms), ncu launch list, one ncu --set full capture
mkdir -p gpurun_out
TAG=${1:-final}
python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests_$TAG.log 2>&1; echo "gpu suite rc=$?"; tail -2 gpurun_out/r02_gputests_$TAG.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02_smoke_$TAG.log
python bench.py > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_reference_arm.err; echo "reference arm rc=$?"
FILTER='regex:^(?!.*(pack_|to_f32_kernel|bias_l2)).*$'
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -k "$FILTER" -c 400 --csv --log-file gpurun_out/r02_launches_$TAG.csv python bench.py --steps 2 --warmup 1 --no-tiled > gpurun_out/ll_ncu.log 2>&1; echo "launch list rc=$?"
# the report itself stays on the box (gpurun_out/ is capped at 64 MiB): only its raw page travels
timeout 1500 ncu --set full --clock-control none -k "$FILTER" -s 340 -c 62 -f -o /tmp/r02_full_$TAG python bench.py --steps 2 --warmup 1 --no-tiled > gpurun_out/full_ncu.log 2>&1; echo "full capture rc=$?"
ncu -i /tmp/r02_full_$TAG.ncu-rep --page raw --csv > gpurun_out/r02_full_${TAG}_raw.csv 2>/dev/null; ls -la /tmp/r02_full_$TAG.ncu-rep gpurun_out/r02_full_${TAG}_raw.csv
