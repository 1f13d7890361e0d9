"""Engineering: graphed forward time of HAT x4 bf16 against the batch (64x64 LR patches): fixed cost per forward."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import HAT_X4  # noqa: E402
from super_resolution_b200 import HAT  # noqa: E402

torch.manual_seed(0)
net = HAT(**HAT_X4).eval().cuda()
net.engine().output_view = True
for B in (16, 64, 128, 256, 64, 16):
    x = torch.rand(B, 3, 64, 64).bfloat16().cuda()
    for _ in range(3):
        net(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        net(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"B={B:3d} tokens={B * 4096:7d} {ms:8.3f} ms  {ms / B:7.3f} ms/patch  launches {net.engine().last_launch_count}")
