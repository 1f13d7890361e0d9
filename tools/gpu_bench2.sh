#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_tile_pipeline.py tests/test_parity_r2.py -m gpu -x -q -k "tile or tiled" 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c5_bench_2gpu.json 2> gpurun_out/c5_bench_2gpu.err; echo "bench2 rc=$?"
tail -2 gpurun_out/c5_bench_2gpu.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c5_bench_2gpu.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'])
print(json.dumps(d['tiled'], indent=1))
PY
