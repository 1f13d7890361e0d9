#!/bin/bash
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/run_configs.py > gpurun_out/r02_configs_8gpu.jsonl 2> gpurun_out/r02_configs_8gpu.err; echo rc=$?
cat gpurun_out/r02_configs_8gpu.jsonl | cut -c1-400
tail -3 gpurun_out/r02_configs_8gpu.err
