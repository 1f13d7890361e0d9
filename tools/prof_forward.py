"""Run a few eager (no CUDA graph) forwards of the bench workload -- the target process for ncu captures.
    ncu --set full --kernel-name regex:gemm_tc --launch-skip 264 --launch-count 3 python tools/prof_forward.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["HATB200_CUDA_GRAPH"] = "0"
import torch  # noqa: E402

from bench import BATCH, HAT_X4, LR  # noqa: E402
from super_resolution_b200 import HAT  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
torch.manual_seed(0)
net = HAT(**HAT_X4).eval().cuda()
x = torch.rand(BATCH, 3, LR, LR).bfloat16().cuda()
for _ in range(n):
    y = net(x)
torch.cuda.synchronize()
print("ok", float(y.float().mean()))
