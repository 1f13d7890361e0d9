"""Host-side driver of one libhatb200 handle: binds a module's state_dict to the C ABI,
owns the scratch workspace (torch caching allocator) and optional CUDA graphs.

PyTorch is plumbing here (device memory, streams); all arithmetic happens in the
hand-written CUDA library.  Nothing in this file computes on tensors.
"""
import ctypes as C
import os

import torch

from . import _lib


def _config_from_module(m) -> _lib.Config:
    c = _lib.Config()
    c.struct_size = C.sizeof(_lib.Config)
    c.in_chans = m.in_chans
    c.embed_dim = m.embed_dim
    c.num_layers = m.num_layers
    if m.num_layers > _lib.MAX_LAYERS:
        raise ValueError(f"at most {_lib.MAX_LAYERS} residual groups are supported")
    for i in range(m.num_layers):
        c.depths[i] = m.depths[i]
        c.num_heads[i] = m.heads[i]
    c.window_size = m.window_size
    c.overlap_win_size = int(m.window_size * m.overlap_ratio) + m.window_size     # hat_arch.py:289
    c.cab_mid = m.embed_dim // m.compress_ratio                                   # hat_arch.py:84
    c.ffn_hidden = int(m.embed_dim * m.mlp_ratio)                                 # hat_arch.py:98
    c.esc_pdim, c.esc_kernel = m.esc_pdim, m.esc_kernel
    c.ocab_esc_enable = int(m.ocab_esc_enable)
    c.ocab_esc_pdim, c.ocab_esc_kernel = m.ocab_esc_pdim, m.ocab_esc_kernel
    c.upscale = m.upscale
    c.upsampler = 1 if m.upsampler == 'pixelshuffle' else 0
    c.resi_conv = 1 if m.resi_connection == '1conv' else 0
    c.patch_norm = int(bool(m.patch_norm))
    c.qkv_bias = int(bool(m.qkv_bias))
    c.conv_scale = float(m.conv_scale)
    c.img_range = float(m.img_range)
    c.qk_scale = float(m.qk_scale) if m.qk_scale else 0.0
    ows = c.overlap_win_size
    hatx = getattr(m, "arch", "HAT") == "HATX"
    c.arch = 1 if hatx else 0
    c.ocab_pad = (ows - m.window_size + 1) // 2 if hatx else (ows - m.window_size) // 2       # hatx_arch.py:316-318
    ratio = float(getattr(m, "kv_topk_ratio", 1.0))
    c.kv_topk_keep = max(1, int(ratio * ows * ows)) if (hatx and ratio < 1.0) else 0         # hatx_arch.py:421
    c.use_focus_bias = int(hatx and bool(getattr(m, "use_focus_bias", False)))
    mean = [float(v) for v in m.mean.flatten().tolist()]
    for i in range(4):
        c.mean[i] = mean[i] if i < len(mean) else (mean[0] if len(mean) == 1 else 0.0)
    return c


class Engine:
    def __init__(self, module):
        self.module = module
        self.lib = _lib.lib()          # raises if the CUDA library was not built: no fallback
        self.handle = None
        self.device = None
        self._bound = None             # tensors borrowed by the handle (kept alive)
        self._sample = None            # staleness probe over every bound tensor
        self._probe_tensors = []
        # pre-packed weight cache: directory of blobs keyed by a hash of (configuration, state_dict);
        # None disables it.  A hit skips every hatb200_set_param and all re-layout kernels of hatb200_pack.
        self.pack_cache_dir = os.environ.get("HATB200_PACK_CACHE") or None
        self.pack_cache_hit = None     # True / False after a (re)bind with the cache enabled
        self.output_view = False       # True: graphed forwards return the graph's own output buffer (valid until the
                                       # next forward of the same shape) instead of a fresh copy of it
        self.output_u8 = False         # True (bf16 inputs): forward returns the image as tensor2img bytes, (B, H*s, W*s, C)
                                       # uint8 BGR written by the conv_last epilogue (hatb200.h, y_dtype HATB200_U8)
        self.last_fallback_count = 0
        self._ws = None
        self._graphs = {}
        self.use_cuda_graph = os.environ.get("HATB200_CUDA_GRAPH", "1") != "0"
        self.options = {}
        self.last_launch_count = 0

    # -- lifetime ---------------------------------------------------------------
    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __deepcopy__(self, memo):
        return None                    # a copied module builds its own engine lazily

    def close(self):
        self._graphs = {}
        if self.handle is not None:
            self.lib.hatb200_destroy(self.handle)
            self.handle = None

    def invalidate(self):
        """Parameters changed (load_state_dict / .to()): re-bind and re-pack at the next forward."""
        self._bound = None
        self._graphs = {}

    def set_option(self, name, value):
        self.options[name] = int(value)
        self._graphs = {}
        if self.handle is not None:
            _lib.check(self.lib.hatb200_set_option(self.handle, name.encode(), int(value)))

    # -- binding ----------------------------------------------------------------
    def _probe(self):
        # staleness check on EVERY parameter / buffer (in-place updates bump _version, .to() / .data = ... move
        # data_ptr): ~0.2 ms of host time for 912 tensors, hidden behind the asynchronous launch of the previous step.
        # load_state_dict / _apply also call invalidate() explicitly
        return tuple((t.data_ptr(), t._version) for t in self._probe_tensors)

    def _ensure_bound(self, device):
        m = self.module
        if self.handle is not None and self.device != device:
            self.close()
            self._bound = None
        if self.handle is None:
            with torch.cuda.device(device):
                h = C.c_void_p()
                cfg = _config_from_module(m)
                _lib.check(self.lib.hatb200_create(C.byref(cfg), C.byref(h)))
                self.handle = h
                self.device = device
                for k, v in self.options.items():
                    _lib.check(self.lib.hatb200_set_option(self.handle, k.encode(), v))
        if self._bound is not None and self._probe() == self._sample:
            return
        sd = m.state_dict(keep_vars=True)
        stream = torch.cuda.current_stream(device)
        cache_file = self._cache_file(sd) if self.pack_cache_dir else None
        if cache_file is not None and os.path.exists(cache_file):
            blob = open(cache_file, "rb").read()
            rc = self.lib.hatb200_pack_import(self.handle, blob, len(blob), C.c_void_p(stream.cuda_stream))
            if rc == 0:
                for k, v in sd.items():
                    if v.device != device:
                        raise RuntimeError(f"parameter '{k}' is on {v.device} but the input is on {device}")
                self.pack_cache_hit = True
                self._bound = {}
                self._probe_tensors = list(sd.values())
                self._sample = self._probe()
                self._graphs = {}
                return
            # a stale or foreign blob: fall through to a normal pack (and overwrite it)
        bound = {}
        for k, v in sd.items():
            t = v.detach()
            if t.device != device:
                raise RuntimeError(f"parameter '{k}' is on {t.device} but the input is on {device}; "
                                   "move the module with .to(device) first")
            if not t.is_contiguous():
                t = t.contiguous()
            shape = (C.c_int64 * max(t.dim(), 1))(*t.shape)
            _lib.check(self.lib.hatb200_set_param(self.handle, k.encode(), C.c_void_p(t.data_ptr()),
                                                  _lib.dtype_code(t.dtype), shape, t.dim()))
            bound[k] = t
        _lib.check(self.lib.hatb200_pack(self.handle, C.c_void_p(stream.cuda_stream)))
        if cache_file is not None:
            self.pack_cache_hit = False
            need = C.c_size_t()
            _lib.check(self.lib.hatb200_packed_bytes(self.handle, C.byref(need)))
            buf = C.create_string_buffer(need.value)
            _lib.check(self.lib.hatb200_packed_export(self.handle, buf, need.value, C.c_void_p(stream.cuda_stream)))
            os.makedirs(self.pack_cache_dir, exist_ok=True)
            tmp = f"{cache_file}.{os.getpid()}.tmp"
            with open(tmp, "wb") as f:
                f.write(buf.raw)
            os.replace(tmp, cache_file)           # atomic: concurrent ranks write identical content
        self._bound = bound
        self._probe_tensors = list(sd.values())
        self._sample = self._probe()
        self._graphs = {}

    def _cache_file(self, sd):
        """<pack_cache_dir>/<sha256 of configuration + every state_dict entry (key, dtype, shape, bytes)>.hbpk"""
        import hashlib
        hsh = hashlib.sha256()
        hsh.update(bytes(_config_from_module(self.module)))
        for k, v in sd.items():
            t = v.detach()
            hsh.update(f"{k}|{t.dtype}|{tuple(t.shape)}|".encode())
            hsh.update(t.contiguous().cpu().view(torch.uint8).numpy().tobytes() if t.numel() else b"")
        return os.path.join(self.pack_cache_dir, hsh.hexdigest() + ".hbpk")

    # -- forward ----------------------------------------------------------------
    def _workspace(self, B, H, W, code, device):
        need = C.c_size_t()
        _lib.check(self.lib.hatb200_workspace_bytes(self.handle, B, H, W, code, C.byref(need)))
        if self._ws is None or self._ws.numel() < need.value or self._ws.device != device:
            self._ws = None
            self._graphs = {}
            self._ws = torch.empty(need.value, dtype=torch.uint8, device=device)
        return self._ws

    def _launch(self, x, y, ws, features=False):
        B, _, H, W = x.shape
        stream = torch.cuda.current_stream(x.device)
        code = _lib.dtype_code(x.dtype)
        fn = self.lib.hatb200_forward_features if features else self.lib.hatb200_forward
        _lib.check(fn(self.handle, C.c_void_p(x.data_ptr()), code, B, H, W,
                      C.c_void_p(y.data_ptr()), _lib.dtype_code(y.dtype),
                      C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(stream.cuda_stream)))
        self.last_launch_count = self.lib.hatb200_last_launch_count(self.handle)
        self.last_fallback_count = self.lib.hatb200_last_fallback_count(self.handle)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        m = self.module
        if not isinstance(x, torch.Tensor) or x.dim() != 4:
            raise RuntimeError("HAT.forward expects a (B, C, H, W) tensor")
        if not x.is_cuda:
            raise RuntimeError("super_resolution_b200.HAT has no CPU path: the input must be a CUDA tensor "
                               "(the forward runs in libhatb200 on sm_100a)")
        if x.shape[1] != m.in_chans:
            raise RuntimeError(f"expected {m.in_chans} input channels, got {x.shape[1]}")
        if x.dtype not in (torch.float32, torch.bfloat16):
            raise NotImplementedError(f"input dtype {x.dtype}: float32 (parity path) and bfloat16 "
                                      "(tensor-core path) are implemented")
        B, _, H, W = x.shape
        ws_ = m.window_size
        if H % ws_ or W % ws_ or H < ws_ or W < ws_:     # reference: RuntimeError from window_partition's .view
            raise RuntimeError(f"shape '[{H}, {W}]' is invalid: height and width must be multiples of "
                               f"window_size {ws_}")
        x = x.contiguous()
        with torch.cuda.device(x.device):
            self._ensure_bound(x.device)
            s = m.upscale if m.upsampler == 'pixelshuffle' else 1
            ws = self._workspace(B, H, W, _lib.dtype_code(x.dtype), x.device)
            if not self.use_cuda_graph or torch.cuda.is_current_stream_capturing():
                y = self._alloc_y(x, s)
                self._launch(x, y, ws)
                return y
            return self._forward_graphed(x, ws, s)

    def _alloc_y(self, x, s):
        B, _, H, W = x.shape
        if self.output_u8:
            if x.dtype != torch.bfloat16:
                raise NotImplementedError("output_u8 is implemented for bfloat16 inputs (the tensor-core path)")
            return torch.empty((B, H * s, W * s, self.module.in_chans), dtype=torch.uint8, device=x.device)
        return torch.empty((B, self.module.in_chans, H * s, W * s), dtype=x.dtype, device=x.device)

    def forward_features(self, x: torch.Tensor) -> torch.Tensor:
        """HAT.forward_features (hat_arch.py:828-846): (B, embed_dim, H, W) -> (B, embed_dim, H, W), eager launches."""
        m = self.module
        if not isinstance(x, torch.Tensor) or x.dim() != 4 or x.shape[1] != m.embed_dim:
            raise RuntimeError(f"HAT.forward_features expects a (B, {m.embed_dim}, H, W) tensor")
        if not x.is_cuda:
            raise RuntimeError("super_resolution_b200.HAT has no CPU path: the input must be a CUDA tensor")
        if x.dtype not in (torch.float32, torch.bfloat16):
            raise NotImplementedError(f"input dtype {x.dtype}: float32 and bfloat16 are implemented")
        B, _, H, W = x.shape
        ws_ = m.window_size
        if H % ws_ or W % ws_ or H < ws_ or W < ws_:
            raise RuntimeError(f"shape '[{H}, {W}]' is invalid: height and width must be multiples of "
                               f"window_size {ws_}")
        x = x.contiguous()
        with torch.cuda.device(x.device):
            self._ensure_bound(x.device)
            ws = self._workspace(B, H, W, _lib.dtype_code(x.dtype), x.device)
            y = torch.empty_like(x)
            self._launch(x, y, ws, features=True)
            return y

    def _forward_graphed(self, x, ws, s):
        key = (tuple(x.shape), x.dtype, bool(self.output_u8))
        entry = self._graphs.get(key)
        if entry is None:
            xs = torch.empty_like(x)
            ys = self._alloc_y(x, s)
            xs.copy_(x)
            self._launch(xs, ys, ws)             # eager warm-up (also surfaces errors outside capture)
            torch.cuda.current_stream().synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._launch(xs, ys, ws)
            if len(self._graphs) >= 8:
                self._graphs.pop(next(iter(self._graphs)))
            entry = self._graphs[key] = (g, xs, ys, ws)
        g, xs, ys, _ = entry
        xs.copy_(x)
        g.replay()
        return ys if self.output_view else ys.clone()

    # -- per-op profile (bench roofline breakdown) ----------------------------------
    def op_profile(self, x: torch.Tensor) -> dict:
        """Run one eager forward with CUDA events around every op and return
        {tag: {kernel, launches, ms, flops, bytes}} (algorithmic flops / bytes per tag, serving kernel)."""
        import json
        graph, self.use_cuda_graph = self.use_cuda_graph, False
        try:
            self.set_option("op_timing", 1)
            self.forward(x)
            buf = C.create_string_buffer(1 << 16)
            _lib.check(self.lib.hatb200_op_profile(self.handle, buf, len(buf)))
            return json.loads(buf.value.decode())
        finally:
            self.set_option("op_timing", 0)
            self.use_cuda_graph = graph
