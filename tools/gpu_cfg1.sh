#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_tile_pipeline.py tests/test_parity_r2.py tests/test_hat_model_r2.py -m gpu -x -q -k "tile or tiled or nondist" 2>&1 | tail -3
python tools/run_configs.py --cpu > gpurun_out/r02_configs_1gpu.jsonl 2> gpurun_out/r02_configs_1gpu.err; echo rc=$?
cat gpurun_out/r02_configs_1gpu.jsonl
