#!/bin/bash
# knob sweep over the fused GEMMs inside the whole forward (tools/time_ops.py)
run() { echo "== $*"; env "$@" python tools/time_ops.py esc_aggr ffn_fc2 ocab_mlp2 ocab_proj 2>&1 | grep -E "us/launch|eager"; }
run HATB200_G3_PRE=0
run HATB200_G3_PRE=1
run HATB200_G3_PRESTAGES=2
run HATB200_G3_PRESTAGES=3
run HATB200_G3_PRESTAGES=4
run HATB200_G3_PRESTAGES=6
run HATB200_G3_PRESTAGES=4 HATB200_G3_OPS=2
run HATB200_G3_PRESTAGES=6 HATB200_G3_OPS=2
run HATB200_G3_PRESTAGES=2 HATB200_G3_STGBUFS=2
run HATB200_G3_PRESTAGES=4 HATB200_G3_STGBUFS=1
run HATB200_G3_PRESTAGES=6 HATB200_G3_STGBUFS=1
