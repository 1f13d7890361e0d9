"""Engineering: graphed HAT-L x4 bf16 forward time against (batch, tile shape) -- the cost model of tile sharding."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import HATL_X4  # noqa: E402
from super_resolution_b200 import HAT  # noqa: E402

torch.manual_seed(0)
net = HAT(**HATL_X4).eval().cuda()
net.engine().output_view = True
for B, h, w in ((1, 288, 288), (2, 288, 288), (1, 288, 320), (2, 288, 320), (1, 320, 320), (2, 320, 320), (4, 320, 320),
                (8, 288, 320), (1, 64, 64), (16, 64, 64)):
    x = torch.rand(B, 3, h, w).bfloat16().cuda()
    for _ in range(2):
        net(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        net(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"B={B} {h}x{w}: {ms:8.2f} ms  {ms * 1e3 / (B * h * w):.4f} us/px")
