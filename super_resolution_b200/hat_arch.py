"""Drop-in `HAT` for the reference class registered in basicsr's ARCH_REGISTRY
(reference: HAT/hat/archs/hat_arch.py:607-859).

Same constructor arguments, same attribute / parameter tree and therefore the same
`state_dict` layout (912 keys for HAT x4), so reference checkpoints load with
`strict=True`.  The modules below only *hold parameters*; the forward is one call into
libhatb200 (hand-written sm_100a CUDA, see csrc/) through the C ABI of
include/hatb200.h.  Forward-only: there is no autograd, no CPU path and no PyTorch
fallback -- `forward` raises if the input is not a CUDA tensor.
"""
import math

import warnings

import torch
import torch.nn as nn

from .engine import Engine

try:  # register under the reference's name when basicsr is importable (hat_arch.py:607)
    from basicsr.utils.registry import ARCH_REGISTRY as _ARCH_REGISTRY
except Exception:  # basicsr is optional: the class works without the registry
    _ARCH_REGISTRY = None


def _trunc_normal_(t, std):
    return nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2.0, b=2.0)


class _Holder(nn.Module):
    """Parameter container; never called."""

    def forward(self, *a, **k):  # pragma: no cover
        raise RuntimeError("parameter holder: the forward runs inside libhatb200")


class _ECA(_Holder):                       # hat_arch.py:66-78
    def __init__(self, k_size=5):
        super().__init__()
        self.conv = nn.Conv1d(1, 1, kernel_size=k_size, padding=k_size // 2, bias=False)


class _CAB(_Holder):                       # hat_arch.py:80-90
    def __init__(self, num_feat, compress_ratio):
        super().__init__()
        mid = num_feat // compress_ratio
        self.cab = nn.Sequential(nn.Conv2d(num_feat, mid, 3, 1, 1), nn.GELU(),
                                 nn.Conv2d(mid, num_feat, 3, 1, 1), _ECA(5))


class _ConvolutionalAttention(_Holder):    # esc_arch.py:89-102
    def __init__(self, pdim):
        super().__init__()
        self.dwc_proj = nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Conv2d(pdim, pdim // 2, 1, 1, 0), nn.GELU(),
                                      nn.Conv2d(pdim // 2, pdim * 9, 1, 1, 0))
        nn.init.zeros_(self.dwc_proj[-1].weight)
        nn.init.zeros_(self.dwc_proj[-1].bias)


class _ConvAttnWrapper(_Holder):           # esc_arch.py:136-145
    def __init__(self, dim, pdim):
        super().__init__()
        self.plk = _ConvolutionalAttention(pdim)
        self.aggr = nn.Conv2d(dim, dim, 1, 1, 0)


class _ESCConvAttn(_Holder):               # hat_arch.py:139-149
    def __init__(self, dim, pdim, kernel_size):
        super().__init__()
        self.core = _ConvAttnWrapper(dim, pdim)
        self.plk_filter = nn.Parameter(torch.randn(pdim, pdim, kernel_size, kernel_size))
        nn.init.orthogonal_(self.plk_filter)


class _GatedDconvFFN(_Holder):             # hat_arch.py:95-105
    def __init__(self, dim, mlp_ratio):
        super().__init__()
        hidden = int(dim * mlp_ratio)
        self.fc1 = nn.Linear(dim, 2 * hidden)
        self.dw = nn.Conv2d(2 * hidden, 2 * hidden, 3, 1, 1, groups=2 * hidden)
        self.fc2 = nn.Linear(hidden, dim)


class _SpatialGateDConvFFN(_Holder):       # HATX, hatx_arch.py:144-158
    def __init__(self, dim, mlp_ratio):
        super().__init__()
        hidden = int(dim * mlp_ratio)
        assert hidden % 2 == 0, f"Hidden({hidden}) must be even for spatial gate split."
        self.fc1 = nn.Linear(dim, hidden)
        self.dw = nn.Conv2d(hidden // 2, hidden // 2, 3, 1, 1, groups=hidden // 2)
        self.fc2 = nn.Linear(hidden, dim)


class _HAB(_Holder):                       # hat_arch.py:172-215; HATX: hatx_arch.py:185-230 (sgfn=True)
    def __init__(self, dim, compress_ratio, mlp_ratio, norm_layer, esc_pdim, esc_kernel, sgfn=False):
        super().__init__()
        self.norm1 = norm_layer(dim)
        self.esc_attn = _ESCConvAttn(dim, esc_pdim, esc_kernel)
        self.conv_block = _CAB(dim, compress_ratio)
        self.norm2 = norm_layer(dim)
        self.mlp = (_SpatialGateDConvFFN if sgfn else _GatedDconvFFN)(dim, mlp_ratio)


class _OCAB(_Holder):                      # hat_arch.py:267-324; HATX: hatx_arch.py:289-365 (focus head)
    def __init__(self, dim, window_size, overlap_ratio, num_heads, qkv_bias, mlp_ratio, norm_layer,
                 esc_enable, esc_pdim, esc_kernel, use_focus_bias=False):
        super().__init__()
        ows = int(window_size * overlap_ratio) + window_size
        self.norm1 = norm_layer(dim)
        self.q_proj = nn.Linear(dim, dim, bias=qkv_bias)
        self.kv_proj = nn.Linear(dim, 2 * dim, bias=qkv_bias)
        self.relative_position_bias_table = nn.Parameter(
            torch.zeros((window_size + ows - 1) * (window_size + ows - 1), num_heads))
        _trunc_normal_(self.relative_position_bias_table, std=.02)
        self.proj = nn.Linear(dim, dim)
        self.norm2 = norm_layer(dim)
        hidden = int(dim * mlp_ratio)
        self.mlp = nn.Sequential(nn.Linear(dim, hidden), nn.GELU(), nn.Linear(hidden, dim))
        if esc_enable:
            self.esc_core = _ConvAttnWrapper(dim, esc_pdim)
            self.esc_plk = nn.Parameter(torch.randn(esc_pdim, esc_pdim, esc_kernel, esc_kernel))
            nn.init.orthogonal_(self.esc_plk)
        if use_focus_bias:
            self.focus_head = nn.Sequential(nn.Conv2d(dim, dim // 4, 1, 1, 0), nn.GELU(), nn.Conv2d(dim // 4, 1, 1, 1, 0))


class _AttenBlocks(_Holder):               # hat_arch.py:395-464
    def __init__(self, dim, depth, num_heads, window_size, compress_ratio, overlap_ratio, mlp_ratio, qkv_bias,
                 norm_layer, esc_pdim, esc_kernel, ocab_esc_enable, ocab_esc_pdim, ocab_esc_kernel,
                 hab_ffn_ratio=None, use_focus_bias=False):
        super().__init__()
        # HATX: HAB uses the spatial-gate FFN.  The reference stores `hab_ffn_ratio` (hatx_arch.py:790) but never passes
        # it on: RHAG / AttenBlocks / HAB all receive `mlp_ratio` (hatx_arch.py:824, :514), so the SGFN hidden width is
        # int(dim * mlp_ratio) -- the reference's behaviour is what the state_dict layout follows.
        sgfn = hab_ffn_ratio is not None
        self.blocks = nn.ModuleList([
            _HAB(dim, compress_ratio, mlp_ratio, norm_layer, esc_pdim, esc_kernel, sgfn) for _ in range(depth)])
        self.overlap_attn = _OCAB(dim, window_size, overlap_ratio, num_heads, qkv_bias, mlp_ratio, norm_layer,
                                  ocab_esc_enable, ocab_esc_pdim, ocab_esc_kernel, use_focus_bias)


class _RHAG(_Holder):                      # hat_arch.py:484-553
    def __init__(self, dim, resi_connection, **kw):
        super().__init__()
        self.residual_group = _AttenBlocks(dim, **kw)
        if resi_connection == '1conv':
            self.conv = nn.Conv2d(dim, dim, 3, 1, 1)
        elif resi_connection == 'identity':
            self.conv = nn.Identity()
        else:
            raise ValueError(f"Unknown resi_connection: {resi_connection}")


class _PatchNorm(_Holder):                 # PatchEmbed with norm_layer, hat_arch.py:558-575
    def __init__(self, dim, norm_layer):
        super().__init__()
        self.norm = norm_layer(dim)


class _Upsample(nn.Sequential):            # hat_arch.py:593-605
    def __init__(self, scale, num_feat):
        m = []
        if (scale & (scale - 1)) == 0:
            for _ in range(int(math.log(scale, 2))):
                m += [nn.Conv2d(num_feat, 4 * num_feat, 3, 1, 1), nn.PixelShuffle(2)]
        elif scale == 3:
            m += [nn.Conv2d(num_feat, 9 * num_feat, 3, 1, 1), nn.PixelShuffle(3)]
        else:
            raise ValueError(f'scale {scale} is not supported. Supported scales: 2^n and 3.')
        super().__init__(*m)


class HAT(nn.Module):
    r"""B200-native HAT (HAT + ESC hybrid of the reference), forward-only.

    Constructor arguments, defaults and ignored arguments follow the reference
    (hat_arch.py:610-644): `squeeze_factor`, `esc_use_dynamic`, `use_checkpoint`,
    `drop_rate`, `attn_drop_rate`, `drop_path_rate` (identity in eval) and unknown
    `**kwargs` are accepted and ignored.  `forward(x)`: `(B, in_chans, H, W)` CUDA tensor,
    H and W multiples of `window_size`; float32 input runs the fp32 parity path
    (max-abs <= 1e-4 vs the reference), bfloat16 input runs the tcgen05 bf16 path
    (PSNR >= 50 dB vs the fp32 reference).  Output dtype = input dtype.
    """

    def __init__(self, img_size=64, patch_size=1, in_chans=3, embed_dim=96, depths=(6, 6, 6, 6),
                 num_heads=(6, 6, 6, 6), window_size=7, compress_ratio=3, squeeze_factor=30, conv_scale=0.01,
                 overlap_ratio=0.5, mlp_ratio=4., qkv_bias=True, qk_scale=None, drop_rate=0., attn_drop_rate=0.,
                 drop_path_rate=0.1, norm_layer=nn.LayerNorm, ape=False, patch_norm=True, use_checkpoint=False,
                 upscale=2, img_range=1., upsampler='', resi_connection='1conv', esc_pdim: int = 16,
                 esc_kernel: int = 13, esc_use_dynamic: bool = True, ocab_esc_enable: bool = False,
                 ocab_esc_pdim: int = 16, ocab_esc_kernel: int = 13, **kwargs):
        super().__init__()
        # HATX (subclass below) passes its three extra arguments through these private keywords
        hab_ffn_ratio = kwargs.pop('_hab_ffn_ratio', None)
        self.kv_topk_ratio = float(kwargs.pop('_kv_topk_ratio', 1.0))
        self.use_focus_bias = bool(kwargs.pop('_use_focus_bias', False))
        self.hab_ffn_ratio = hab_ffn_ratio
        self.arch = 'HATX' if hab_ffn_ratio is not None else 'HAT'
        if ape:
            raise NotImplementedError("ape=True (absolute position embedding, hat_arch.py:699-702) is not "
                                      "implemented; no shipped configuration uses it")
        if norm_layer is not nn.LayerNorm:
            raise NotImplementedError("only norm_layer=nn.LayerNorm is implemented")
        if len(depths) != len(num_heads):
            raise ValueError("depths and num_heads must have the same length")
        self.window_size = window_size
        self.shift_size = window_size // 2
        self.overlap_ratio = overlap_ratio
        num_feat = 64
        self.img_range = img_range
        if in_chans == 3:
            self.mean = torch.Tensor((0.4488, 0.4371, 0.4040)).view(1, 3, 1, 1)
        else:
            self.mean = torch.zeros(1, 1, 1, 1)
        self.upscale = upscale
        self.upsampler = upsampler
        self.in_chans = in_chans
        self.conv_scale = conv_scale
        self.compress_ratio = compress_ratio
        self.qk_scale = qk_scale
        self.qkv_bias = qkv_bias
        self.depths = tuple(depths)
        self.heads = tuple(num_heads)
        self.resi_connection = resi_connection
        self.esc_pdim, self.esc_kernel = esc_pdim, esc_kernel
        self.ocab_esc_enable, self.ocab_esc_pdim, self.ocab_esc_kernel = bool(ocab_esc_enable), ocab_esc_pdim, ocab_esc_kernel

        self.register_buffer('relative_position_index_SA', self.calculate_rpi_sa())
        self.register_buffer('relative_position_index_OCA', self.calculate_rpi_oca())

        self.conv_first = nn.Conv2d(in_chans, embed_dim, 3, 1, 1)
        self.num_layers = len(depths)
        self.embed_dim = embed_dim
        self.ape = ape
        self.patch_norm = patch_norm
        self.num_features = embed_dim
        self.mlp_ratio = mlp_ratio
        isz = (img_size, img_size) if isinstance(img_size, int) else tuple(img_size)
        psz = (patch_size, patch_size) if isinstance(patch_size, int) else tuple(patch_size)
        self.patches_resolution = [isz[0] // psz[0], isz[1] // psz[1]]
        self.patch_embed = _PatchNorm(embed_dim, norm_layer) if patch_norm else _Holder()

        self.layers = nn.ModuleList([
            _RHAG(embed_dim, resi_connection, depth=depths[i], num_heads=num_heads[i], window_size=window_size,
                  compress_ratio=compress_ratio, overlap_ratio=overlap_ratio, mlp_ratio=mlp_ratio, qkv_bias=qkv_bias,
                  norm_layer=norm_layer, esc_pdim=esc_pdim, esc_kernel=esc_kernel, ocab_esc_enable=ocab_esc_enable,
                  ocab_esc_pdim=ocab_esc_pdim, ocab_esc_kernel=ocab_esc_kernel, hab_ffn_ratio=hab_ffn_ratio,
                  use_focus_bias=self.use_focus_bias)
            for i in range(self.num_layers)])
        self.norm = norm_layer(self.num_features)

        if resi_connection == '1conv':
            self.conv_after_body = nn.Conv2d(embed_dim, embed_dim, 3, 1, 1)
        elif resi_connection == 'identity':
            self.conv_after_body = nn.Identity()
        else:
            raise ValueError(f"Unknown resi_connection: {resi_connection}")

        if self.upsampler == 'pixelshuffle':
            self.conv_before_upsample = nn.Sequential(nn.Conv2d(embed_dim, num_feat, 3, 1, 1),
                                                      nn.LeakyReLU(inplace=True))
            self.upsample = _Upsample(upscale, num_feat)
            self.conv_last = nn.Conv2d(num_feat, in_chans, 3, 1, 1)

        self.apply(self._init_weights)
        self._engine = None
        self._warned_grad = False

    def _init_weights(self, m):              # hat_arch.py:761-768
        if isinstance(m, nn.Linear):
            _trunc_normal_(m.weight, std=.02)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.LayerNorm):
            nn.init.constant_(m.bias, 0)
            nn.init.constant_(m.weight, 1.0)

    # ---- index tables (integer bookkeeping, bit-exact with the reference) ----
    def calculate_rpi_sa(self):              # hat_arch.py:770-781
        ws = self.window_size
        pos = torch.arange(ws * ws)
        y, x = pos // ws, pos % ws
        return (y[:, None] - y[None, :] + ws - 1) * (2 * ws - 1) + (x[:, None] - x[None, :] + ws - 1)

    def calculate_rpi_oca(self):             # hat_arch.py:783-803 (negative entries are intended)
        ws = self.window_size
        wse = ws + int(self.overlap_ratio * ws)
        q = torch.arange(ws * ws)
        k = torch.arange(wse * wse)
        off = ws - wse + 1
        dy = (k // wse)[None, :] - (q // ws)[:, None] + off
        dx = (k % wse)[None, :] - (q % ws)[:, None] + off
        return dy * (ws + wse - 1) + dx

    def calculate_mask(self, x_size):        # hat_arch.py:805-818 (API compatibility; never used by forward)
        h, w = x_size
        ws, sh = self.window_size, self.shift_size
        if h % ws or w % ws:
            raise RuntimeError(f"calculate_mask: {h}x{w} is not a multiple of window_size {ws}")
        region = torch.zeros(h, w)
        bounds = lambda n: ((0, n - ws), (n - ws, n - sh), (n - sh, n))
        cnt = 0
        for y0, y1 in bounds(h):
            for x0, x1 in bounds(w):
                region[y0:y1, x0:x1] = cnt
                cnt += 1
        win = region.view(h // ws, ws, w // ws, ws).permute(0, 2, 1, 3).reshape(-1, ws * ws)
        diff = win[:, None, :] - win[:, :, None]
        return torch.where(diff != 0, torch.full_like(diff, -100.0), torch.zeros_like(diff))

    @torch.jit.ignore
    def no_weight_decay(self):
        return {'absolute_pos_embed'}

    @torch.jit.ignore
    def no_weight_decay_keywords(self):
        return {'relative_position_bias_table'}

    # ---- compute -------------------------------------------------------------
    def engine(self) -> Engine:
        if self._engine is None:
            self._engine = Engine(self)
        return self._engine

    def _forward_only_guard(self, x):
        if torch.is_grad_enabled():
            # forward-only build: never hand back a graph-less tensor silently
            if x.requires_grad:
                raise RuntimeError("super_resolution_b200.HAT is forward-only: call it under torch.no_grad()")
            if not self._warned_grad and any(p.requires_grad for p in self.parameters()):
                self._warned_grad = True
                warnings.warn("super_resolution_b200.HAT is forward-only: parameters require grad and grad mode is "
                              "enabled, but the output carries no autograd graph (call under torch.no_grad())",
                              RuntimeWarning, stacklevel=3)
        if self.training:
            raise RuntimeError("super_resolution_b200.HAT is inference-only: call .eval() first "
                               "(train mode of the reference uses stochastic depth, hat_arch.py:46-54)")

    def forward_features(self, x):           # hat_arch.py:828-846: (B, embed_dim, H, W) -> (B, embed_dim, H, W)
        self._forward_only_guard(x)
        with torch.no_grad():
            return self.engine().forward_features(x)

    def forward(self, x):                    # hat_arch.py:848-859
        self._forward_only_guard(x)
        self.mean = self.mean.type_as(x)
        with torch.no_grad():
            return self.engine().forward(x)

    def _apply(self, fn, *a, **k):           # .to() / .cuda() / .bfloat16(): packed weights are stale
        out = super()._apply(fn, *a, **k)
        if getattr(self, "_engine", None) is not None:
            self._engine.invalidate()
        return out

    def load_state_dict(self, *a, **k):
        out = super().load_state_dict(*a, **k)
        if getattr(self, "_engine", None) is not None:
            self._engine.invalidate()
        return out


class HATX(HAT):
    r"""B200-native `HATX`, the sibling variant of the reference (HAT/hat/archs/hatx_arch.py:708-975): HAT whose HAB
    uses the Spatial-Gate DConv FFN (hidden width int(dim * mlp_ratio): `hab_ffn_ratio` is accepted and, as in the
    reference, unused; fc1 -> split -> depthwise 3x3 on the first half -> cat[xa *
    SiLU(xb), xb] -> fc2, hatx_arch.py:144-180) and whose OCAB pads its overlap window by the ceiling
    (hatx_arch.py:316-322) and optionally adds a per-key focus bias and prunes keys to the top `kv_topk_ratio`
    (hatx_arch.py:405-441).  Same state_dict layout as the reference class.  With the options at their defaults the
    OCAB runs on the tcgen05 attention kernel; with focus bias / top-k / an odd overlap it runs on the CUDA-core
    attention (counted by `engine().last_fallback_count`).  Ties of the top-k at the threshold (only possible among
    zero-padded keys) are broken by the lower key index; the reference's torch.topk leaves them unspecified."""

    def __init__(self, *args, hab_ffn_ratio: float = 2.0, kv_topk_ratio: float = 1.0, use_focus_bias: bool = False,
                 **kwargs):
        super().__init__(*args, _hab_ffn_ratio=float(hab_ffn_ratio), _kv_topk_ratio=kv_topk_ratio,
                         _use_focus_bias=use_focus_bias, **kwargs)


if _ARCH_REGISTRY is not None:
    try:
        _ARCH_REGISTRY.register(HATX)
    except AssertionError:
        _ARCH_REGISTRY.register(HATX, suffix='b200')
    try:
        _ARCH_REGISTRY.register(HAT)
    except AssertionError:
        # the reference's own hat_arch.HAT is already registered in this process (tests that
        # hold both): keep ours importable under a suffix instead of failing the import
        _ARCH_REGISTRY.register(HAT, suffix='b200')
