"""Per-CTA timeline of one persistent GEMM launch (engineering aid, HATB200_GEMM_TRACE=1)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["HATB200_GEMM_TRACE"] = "1"
import numpy as np, torch
from super_resolution_b200 import _lib
from tests.test_gpu_kernels import _p, _stream  # noqa
lib = C.CDLL(_lib.LIB_PATH)
L = _lib.lib()
M_B, H, W = 16, 64, 64
Cin, N = int(sys.argv[1]) if len(sys.argv) > 1 else 180, int(sys.argv[2]) if len(sys.argv) > 2 else 720
pitch = ((Cin + 63) // 64) * 64
x = torch.zeros(M_B * H * W, pitch, dtype=torch.bfloat16, device="cuda"); x[:, :Cin] = torch.randn(M_B * H * W, Cin, device="cuda").bfloat16()
w = torch.randn(N, Cin, 1, 1, device="cuda") * 0.05
b = torch.randn(N, device="cuda")
opitch = ((N + 63) // 64) * 64
out = torch.zeros(M_B * H * W, opitch, dtype=torch.bfloat16, device="cuda")
for _ in range(3):
    _lib.check(L.hatb200_conv_tokens(_p(x), _lib.BF16, pitch, 0, Cin, M_B, H, W, _p(w), _p(b), N, 1, 1, 0, _p(out), opitch, 1, _stream()))
torch.cuda.synchronize()
buf = np.zeros((1024, 64), dtype=np.uint64)
lib.hatb200_debug_gemm_trace.argtypes = [C.c_void_p, C.c_void_p]
rc = lib.hatb200_debug_gemm_trace(buf.ctypes.data, None)
t = buf.astype(np.int64)
live = t[:, 0] > 0
t0 = t[live, 0].min()
print("ctas", live.sum())
for cta in list(np.nonzero(live)[0][:2]) + [int(np.nonzero(live)[0][-1])]:
    r = t[cta]
    ev = {k: int(r[k] - t0) for k in range(64) if r[k] > 0}
    print("cta", cta, {k: round(v / 1000, 2) for k, v in sorted(ev.items())})
