#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_kernels2.py -m gpu -x -q -k "ffn_tail" 2>&1 | tail -2
python tools/trace_ffn.py 2>&1 | tail -5
HATB200_FFN_DBG=8 python tools/trace_ffn.py 2>&1 | tail -5
