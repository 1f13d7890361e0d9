#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | grep -v "^$" > gpurun_out/c4_gputests.log; echo "gpu suite rc=${PIPESTATUS[0]}"
tail -4 gpurun_out/c4_gputests.log
python bench.py --no-cpu-baseline > gpurun_out/c4_bench.json 2> gpurun_out/c4_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c4_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'])
for k,v in d['per_op'].items(): print(k, v)
PY
