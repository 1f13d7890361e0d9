"""Engineering: per-op device times of one eager forward of the bench workload (HAT x4, 16 x 64 x 64, bf16).
    python tools/time_ops.py [tag substring ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import HAT_X4, BATCH, LR  # noqa: E402
from super_resolution_b200 import HAT  # noqa: E402

torch.manual_seed(0)
net = HAT(**HAT_X4).eval().cuda()
x = torch.rand(BATCH, 3, LR, LR).bfloat16().cuda()
for _ in range(2):
    net(x)
eng = net.engine()
prof = eng.op_profile(x)
prof = eng.op_profile(x)
tot = sum(r["ms"] for r in prof.values())
keys = sys.argv[1:]
for t, r in prof.items():
    if not keys or any(k in t for k in keys):
        print(f"{t:22s} {r['kernel']:14s} n={r['launches']:3d} {r['ms']:.4f} ms  {r['ms'] * 1e3 / r['launches']:.1f} us/launch")
print(f"eager sum {tot:.3f} ms", os.environ.get("HATB200_FFN_DBG", ""))
