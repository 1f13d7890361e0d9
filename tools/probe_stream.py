# NOTE: needs a library built with the engineering probes: HATB200_ENGINEERING=1 python super_resolution_b200/build.py --force
"""What the memory system gives the esc_aggr traffic pattern (2 bf16 + 1 fp32 streams in, 1 fp32 + 1 bf16 out, 170 MB per
launch at M = 65536 x 192) with no arithmetic and no tensor core: the practical HBM ceiling for a ~30 us kernel."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from super_resolution_b200 import _lib
lib = C.CDLL(_lib.LIB_PATH)
lib.hatb200_probe_stream.argtypes = [C.c_void_p] * 5 + [C.c_size_t, C.c_int, C.c_int, C.c_void_p]
M, N = 65536, 192
n8 = M * N // 8
sets = []
for _ in range(6):
    sets.append((torch.randn(M, N, device="cuda").bfloat16(), torch.randn(M, N, device="cuda").bfloat16(),
                 torch.randn(M, N, device="cuda"), torch.empty(M, N, device="cuda"),
                 torch.empty(M, N, dtype=torch.bfloat16, device="cuda")))
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
bytes_ = M * N * (2 + 2 + 4 + 4 + 2)
for ctas in (148, 296, 592, 1184):
    def go(k):
        a, u, r, o32, o16 = sets[k % len(sets)]
        lib.hatb200_probe_stream(a.data_ptr(), u.data_ptr(), r.data_ptr(), o32.data_ptr(), o16.data_ptr(), n8, ctas, 0, st)
    for k in range(6):
        go(k)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(60):
        go(k)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1000 / 60
    print(f"ctas {ctas}: {us:.1f} us per launch, {bytes_ / us / 1e3:.0f} GB/s (rotating over {len(sets)} buffer sets, back to back)")

# producer -> consumer chains: launch k reads the fp32 and bf16 outputs of launch k - 1 (75 % of its reads) plus one cold
# bf16 stream; same direction every launch vs alternating directions (the consumer starts where the producer ended)
for mode in ("same direction", "alternating"):
    for ctas in (148, 592):
        def go(k):
            pa, pu, pr, po32, po16 = sets[(k - 1) % len(sets)]
            a, u, r, o32, o16 = sets[k % len(sets)]
            rev = (k & 1) if mode == "alternating" else 0
            lib.hatb200_probe_stream(po16.data_ptr(), u.data_ptr(), po32.data_ptr(), o32.data_ptr(), o16.data_ptr(), n8, ctas, rev, st)
        for k in range(6):
            go(k)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(60):
            go(k)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1000 / 60
        print(f"chain, {mode}, ctas {ctas}: {us:.1f} us per launch ({bytes_ / us / 1e3:.0f} GB/s algorithmic)")
