#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s 2>&1 | grep -v "^$" > gpurun_out/c2_gputests.log; echo "gpu suite rc=${PIPESTATUS[0]}"
grep -E "r2_|passed|failed|Error|error" gpurun_out/c2_gputests.log | tail -40
python bench.py > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err; echo "bench rc=$?"
