#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_forward.py tests/test_parity_r2.py tests/test_postproc.py tests/test_hat_model_r2.py -m gpu -q -x > gpurun_out/last_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/last_tests.log
python tools/time_ops.py conv_last upsample 2>&1 | grep -E "us/launch|eager"
HATB200_CONV_LAST_2PASS=1 python tools/time_ops.py conv_last 2>&1 | grep -E "us/launch"
