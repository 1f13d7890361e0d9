#!/bin/bash
python -m pytest tests/test_hat_model_r2.py tests/test_tile_pipeline.py tests/test_gpu_forward.py -m gpu -x -q 2>&1 | tail -5
