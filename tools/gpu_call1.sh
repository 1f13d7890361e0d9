#!/bin/bash
# round-2 first GPU call: no-alloc entry alone + whole GPU suite + baseline bench + sanitizer passes
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/c1_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
python -m pytest tests -m gpu -x -q > gpurun_out/c1_gputests.log 2>&1; echo "gpu suite rc=$?"
tail -3 gpurun_out/c1_gputests.log
python bench.py > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo "bench rc=$?"
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/c1_memcheck.log 2>&1; echo "memcheck rc=$?"
tail -5 gpurun_out/c1_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --print-limit 20 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "full_size or attention_tcgen05 or dwconv_gate or layernorm" > gpurun_out/c1_racecheck.log 2>&1; echo "racecheck rc=$?"
tail -5 gpurun_out/c1_racecheck.log
