#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | grep -v "^$" > gpurun_out/c3_gputests.log; echo "gpu suite rc=${PIPESTATUS[0]}"
tail -30 gpurun_out/c3_gputests.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
