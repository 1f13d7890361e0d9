# NOTE: needs a library built with the engineering probes: HATB200_ENGINEERING=1 python super_resolution_b200/build.py --force
"""tcgen05.cp.128x256b probe: prints, per mode, whether TMEM[row][col] == src[row][col], else the observed mapping."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from super_resolution_b200 import _lib
lib = C.CDLL(_lib.LIB_PATH)
lib.hatb200_probe_tmem_cp.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
src = (torch.arange(128 * 32, dtype=torch.float32, device="cuda")).view(128, 32).contiguous()   # value = row * 32 + col
for mode in (0, 1):
    out = torch.full((128, 32), -1.0, device="cuda")
    rc = lib.hatb200_probe_tmem_cp(src.data_ptr(), mode, out.data_ptr(), None)
    torch.cuda.synchronize()
    ok = torch.equal(out, src)
    print("mode", mode, "rc", rc, "identity" if ok else "NOT identity")
    if not ok:
        o = out.long().cpu()
        for r in (0, 1, 2, 7, 8, 9, 31, 32, 64, 127):
            print("  row", r, "-> src (row, col) of its first 12 columns:", [(int(v) // 32, int(v) % 32) for v in o[r, :12]])
