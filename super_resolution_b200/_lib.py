"""ctypes binding of libhatb200.so (the C ABI in include/hatb200.h).

The library is built in-tree by `super_resolution_b200/build.py` (nvcc, sm_100a).  There is
no fallback of any kind: if the shared object is missing or fails to load, importing the
compute path raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhatb200.so")
MAX_LAYERS = 32
NOALLOC_MAX_HID = 1024      # HATB200_NOALLOC_MAX_HID

F32, BF16, F16, I64, U8 = 0, 1, 2, 3, 4


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("in_chans", C.c_int32), ("embed_dim", C.c_int32), ("num_layers", C.c_int32),
        ("depths", C.c_int32 * MAX_LAYERS), ("num_heads", C.c_int32 * MAX_LAYERS),
        ("window_size", C.c_int32), ("overlap_win_size", C.c_int32), ("cab_mid", C.c_int32),
        ("ffn_hidden", C.c_int32), ("esc_pdim", C.c_int32), ("esc_kernel", C.c_int32),
        ("ocab_esc_enable", C.c_int32), ("ocab_esc_pdim", C.c_int32), ("ocab_esc_kernel", C.c_int32),
        ("upscale", C.c_int32), ("upsampler", C.c_int32), ("resi_conv", C.c_int32), ("patch_norm", C.c_int32),
        ("qkv_bias", C.c_int32), ("conv_scale", C.c_float), ("img_range", C.c_float), ("qk_scale", C.c_float),
        ("mean", C.c_float * 4),
        ("arch", C.c_int32), ("ocab_pad", C.c_int32), ("kv_topk_keep", C.c_int32), ("use_focus_bias", C.c_int32),
    ]


class HatB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libhatb200 error {code}: {msg}")
        self.code = code


_lib = None

_SIGNATURES = {
    "hatb200_version": (C.c_int, []),
    "hatb200_last_error": (C.c_char_p, []),
    "hatb200_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "hatb200_destroy": (None, [C.c_void_p]),
    "hatb200_set_param": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.c_int]),
    "hatb200_num_params": (C.c_int, [C.c_void_p]),
    "hatb200_param_key": (C.c_char_p, [C.c_void_p, C.c_int]),
    "hatb200_pack": (C.c_int, [C.c_void_p, C.c_void_p]),
    "hatb200_packed_bytes": (C.c_int, [C.c_void_p, C.POINTER(C.c_size_t)]),
    "hatb200_packed_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "hatb200_pack_import": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "hatb200_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int]),
    "hatb200_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "hatb200_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                  C.c_void_p, C.c_size_t, C.c_void_p]),
    "hatb200_forward_features": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                           C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "hatb200_last_launch_count": (C.c_int, [C.c_void_p]),
    "hatb200_last_fallback_count": (C.c_int, [C.c_void_p]),
    "hatb200_op_profile": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "hatb200_tile_plan": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int32), C.c_int]),
    "hatb200_crop_place": (C.c_int, [C.c_void_p, C.c_int] + [C.c_int] * 6 + [C.c_void_p] + [C.c_int] * 6 + [C.c_void_p]),
    "hatb200_reflect_pad": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                      C.c_int, C.c_void_p]),
    "hatb200_tensor2img": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int,
                                     C.c_void_p, C.c_void_p]),
    "hatb200_psnr_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                  C.c_void_p, C.c_void_p]),
    "hatb200_ssim_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                  C.c_void_p, C.c_void_p]),
    "hatb200_layernorm": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                    C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    "hatb200_conv_tokens": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                      C.c_int, C.c_void_p]),
    "hatb200_linear_fused": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_float,
                                       C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "hatb200_conv_fused": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p]),
    "hatb200_ocab_attention": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "hatb200_dwconv_gate": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "hatb200_esc_plk": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                  C.c_void_p]),
    "hatb200_conv_first": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]),
    "hatb200_conv_last": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                    C.c_void_p, C.c_float, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "hatb200_hab_finish": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "hatb200_conv_shuffle": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                       C.c_void_p, C.c_void_p]),
    "hatb200_ffn_tail_fused": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_int, C.c_void_p]),
    "hatb200_dwconv_gate_noalloc": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
}


def exported_symbols():
    return sorted(_SIGNATURES)


def lib():
    """Load libhatb200.so (once).  Raises if it was not built -- there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python super_resolution_b200/build.py` "
                "(nvcc, sm_100a).  This package has no CPU or PyTorch fallback.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        if l.hatb200_version() != 101:
            raise ImportError(f"libhatb200 version {l.hatb200_version()} does not match this binding (101)")
        _lib = l
    return _lib


def check(code):
    if code != 0:
        raise HatB200Error(code, lib().hatb200_last_error().decode("utf-8", "replace"))


def dtype_code(torch_dtype):
    import torch
    try:
        return {torch.float32: F32, torch.bfloat16: BF16, torch.float16: F16, torch.int64: I64, torch.uint8: U8}[torch_dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {torch_dtype}") from None


def tile_plan(height, width, tile_size, tile_pad, scale):
    """HATModel.tile_process geometry (hat_model.py:52-108) from the C library -> list of dicts."""
    l = lib()
    n = l.hatb200_tile_plan(height, width, tile_size, tile_pad, scale, None, 0)
    if n < 0:
        check(n)
    buf = (C.c_int32 * (12 * n))()
    l.hatb200_tile_plan(height, width, tile_size, tile_pad, scale, buf, n)
    names = ("in_x0", "in_x1", "in_y0", "in_y1", "out_x0", "out_x1", "out_y0", "out_y1",
             "crop_x0", "crop_x1", "crop_y0", "crop_y1")
    return [dict(idx=i, **{k: int(buf[12 * i + j]) for j, k in enumerate(names)}) for i in range(n)]
