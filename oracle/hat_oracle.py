"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference HAT forward.

This file is the parity *checker*.  It is imported only by `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py`.  The product (`super_resolution_b200/`) never imports it and has no
CPU path at all.

It restates, as plain functions over a `state_dict`, the eval-mode forward of the
reference class `HAT` (reference: HAT/hat/archs/hat_arch.py:607-859 and
HAT/hat/archs/esc_arch.py:89-145).  Each function cites the lines it follows.
Pinning: `tests/test_oracle_golden.py` checks it against fixtures produced by the
*unmodified* reference run in the build container (`oracle/make_golden.py`), and,
when `/root/reference` is mounted, against the live reference.

Deliberate differences from the reference (results identical):
  * `calculate_mask` is not evaluated in the forward -- its result is never read
    by any block (hat_arch.py:217, :831); the divisibility check it implied is
    explicit here.
  * batches are handled per sample, which is what the reference's eval branch
    computes for B=1 and what its train branch computes for B>1
    (esc_arch.py:105-123).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

RGB_MEAN = (0.4488, 0.4371, 0.4040)  # hat_arch.py:659


# --------------------------------------------------------------------------
# integer bookkeeping (bit-exact targets)
# --------------------------------------------------------------------------
def rpi_oca(window_size: int, overlap_ratio: float) -> np.ndarray:
    """relative_position_index_OCA, raw (may be negative). hat_arch.py:783-803."""
    ws = window_size
    wse = ws + int(overlap_ratio * ws)
    qy, qx = np.divmod(np.arange(ws * ws, dtype=np.int64), ws)
    ky, kx = np.divmod(np.arange(wse * wse, dtype=np.int64), wse)
    off = ws - wse + 1
    dy = ky[None, :] - qy[:, None] + off
    dx = kx[None, :] - qx[:, None] + off
    return dy * (ws + wse - 1) + dx


def rpi_sa(window_size: int) -> np.ndarray:
    """relative_position_index_SA. hat_arch.py:770-781."""
    ws = window_size
    y, x = np.divmod(np.arange(ws * ws, dtype=np.int64), ws)
    dy = y[:, None] - y[None, :] + ws - 1
    dx = x[:, None] - x[None, :] + ws - 1
    return dy * (2 * ws - 1) + dx


def wrap_index(idx: np.ndarray, table_rows: int) -> np.ndarray:
    """PyTorch negative-index semantics of `table[idx]` (hat_arch.py:378)."""
    return np.where(idx < 0, idx + table_rows, idx)


def window_token_index(h: int, w: int, ws: int) -> np.ndarray:
    """(nW, ws*ws) flat token index y*w+x of each query in window order.
    window_partition, hat_arch.py:124-128."""
    idx = np.arange(h * w, dtype=np.int64).reshape(h // ws, ws, w // ws, ws)
    return idx.transpose(0, 2, 1, 3).reshape(-1, ws * ws)


def overlap_window_token_index(h: int, w: int, ws: int, ows: int) -> np.ndarray:
    """(nW, ows*ows) flat token index of each key/value of each overlapping
    window, -1 where nn.Unfold zero-pads. hat_arch.py:296-297, :359-364."""
    pad = (ows - ws) // 2
    out = np.full((h // ws, w // ws, ows, ows), -1, dtype=np.int64)
    for wy in range(h // ws):
        for wx in range(w // ws):
            ys = wy * ws - pad + np.arange(ows)
            xs = wx * ws - pad + np.arange(ows)
            ok = ((ys >= 0) & (ys < h))[:, None] & ((xs >= 0) & (xs < w))[None, :]
            flat = ys[:, None] * w + xs[None, :]
            out[wy, wx] = np.where(ok, flat, -1)
    return out.reshape(-1, ows * ows)


def pixel_shuffle_source(c_out: int, h: int, w: int, r: int) -> np.ndarray:
    """For out[c, y, x] of PixelShuffle(r): flat source index into in[(c*r*r), h, w].
    out[c, h*r+i, w*r+j] = in[c*r*r + i*r + j, h, w]."""
    c, y, x = np.meshgrid(np.arange(c_out), np.arange(h * r), np.arange(w * r), indexing="ij")
    cin = c * r * r + (y % r) * r + (x % r)
    return (cin * h + y // r) * w + x // r


def tile_rects(height: int, width: int, tile_size: int, tile_pad: int, scale: int):
    """Tile geometry of HATModel.tile_process (hat_model.py:52-108), in the
    reference's tile order.  Returns a list of dicts of plain ints."""
    tiles_x = math.ceil(width / tile_size)
    tiles_y = math.ceil(height / tile_size)
    rects = []
    for y in range(tiles_y):
        for x in range(tiles_x):
            x0, y0 = x * tile_size, y * tile_size
            x1, y1 = min(x0 + tile_size, width), min(y0 + tile_size, height)
            x0p, x1p = max(x0 - tile_pad, 0), min(x1 + tile_pad, width)
            y0p, y1p = max(y0 - tile_pad, 0), min(y1 + tile_pad, height)
            rects.append(dict(
                idx=y * tiles_x + x,
                in_x0=x0p, in_x1=x1p, in_y0=y0p, in_y1=y1p,
                out_x0=x0 * scale, out_x1=x1 * scale, out_y0=y0 * scale, out_y1=y1 * scale,
                crop_x0=(x0 - x0p) * scale, crop_x1=(x0 - x0p) * scale + (x1 - x0) * scale,
                crop_y0=(y0 - y0p) * scale, crop_y1=(y0 - y0p) * scale + (y1 - y0) * scale))
    return rects


def mod_pad(h: int, w: int, window_size: int):
    """HATModel.pre_process pad amounts (hat_model.py:16-26)."""
    ph = (window_size - h % window_size) % window_size
    pw = (window_size - w % window_size) % window_size
    return ph, pw


# --------------------------------------------------------------------------
# floating-point forward (per-module restatement)
# --------------------------------------------------------------------------
def default_cfg(**over):
    """Constructor defaults of the reference HAT (hat_arch.py:610-644)."""
    cfg = dict(img_size=64, patch_size=1, in_chans=3, embed_dim=96, depths=(6, 6, 6, 6),
               num_heads=(6, 6, 6, 6), window_size=7, compress_ratio=3, squeeze_factor=30,
               conv_scale=0.01, overlap_ratio=0.5, mlp_ratio=4., qkv_bias=True, qk_scale=None,
               upscale=2, img_range=1., upsampler='', resi_connection='1conv',
               esc_pdim=16, esc_kernel=13, ocab_esc_enable=False, ocab_esc_pdim=16,
               ocab_esc_kernel=13, patch_norm=True, ape=False)
    cfg.update(over)
    return cfg


def _ln(x, sd, prefix, eps=1e-5):
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + ".weight"], sd[prefix + ".bias"], eps)


def _conv(x, sd, prefix, padding):
    return F.conv2d(x, sd[prefix + ".weight"], sd.get(prefix + ".bias"), padding=padding)


def _to_img(x, h, w):
    b, n, c = x.shape
    return x.transpose(1, 2).reshape(b, c, h, w)


def _to_tok(x):
    return x.flatten(2).transpose(1, 2)


def eca(x, sd, prefix):
    """ECA.forward, hat_arch.py:72-78: GAP -> Conv1d(1,1,5,pad 2, no bias) over the
    channel axis -> sigmoid -> rescale."""
    y = x.mean(dim=(2, 3))                                   # (B, C)
    y = F.conv1d(y[:, None, :], sd[prefix + ".conv.weight"], padding=2)[:, 0]
    return x * torch.sigmoid(y)[:, :, None, None]


def cab(x, sd, prefix):
    """CAB.forward, hat_arch.py:83-90: conv3x3 -> exact GELU -> conv3x3 -> ECA."""
    t = F.gelu(_conv(x, sd, prefix + ".cab.0", 1))
    u = _conv(t, sd, prefix + ".cab.2", 1)
    return eca(u, sd, prefix + ".cab.3")


def conv_attn(x, sd, core_prefix, plk_filter, pdim):
    """ConvAttnWrapper.forward (esc_arch.py:142-145) over ConvolutionalAttention's
    eval branch (esc_arch.py:119-123), one dynamic kernel set per sample.  The
    reference mutates a *copy* of its input; here the input is left untouched."""
    outs = []
    for i in range(x.shape[0]):
        xi = x[i:i + 1]
        x1 = xi[:, :pdim]
        g = x1.mean(dim=(2, 3), keepdim=True)                # AdaptiveAvgPool2d(1), esc_arch.py:96
        g = F.gelu(_conv(g, sd, core_prefix + ".plk.dwc_proj.1", 0))
        dyn = _conv(g, sd, core_prefix + ".plk.dwc_proj.3", 0).reshape(pdim, 1, 3, 3)
        k = plk_filter.shape[-1]
        y1 = F.conv2d(x1, plk_filter, padding=k // 2) + F.conv2d(x1, dyn, padding=1, groups=pdim)
        outs.append(torch.cat([y1, xi[:, pdim:]], dim=1))
    y = torch.cat(outs, dim=0)
    return _conv(y, sd, core_prefix + ".aggr", 0)


def gated_dconv_ffn(x, sd, prefix, h, w):
    """GatedDconvFFN.forward, hat_arch.py:107-119."""
    t = F.linear(x, sd[prefix + ".fc1.weight"], sd[prefix + ".fc1.bias"])
    t = _to_img(t, h, w)
    t = F.conv2d(t, sd[prefix + ".dw.weight"], sd[prefix + ".dw.bias"], padding=1, groups=t.shape[1])
    t = _to_tok(t)
    proj, gate = t.chunk(2, dim=-1)
    t = proj * F.silu(gate)
    return F.linear(t, sd[prefix + ".fc2.weight"], sd[prefix + ".fc2.bias"])


def spatial_gate_ffn(x, sd, prefix, h, w):
    """SpatialGateDConvFFN.forward of the sibling HATX, hatx_arch.py:160-180: fc1 -> split in halves (xa | xb) ->
    depthwise 3x3 on xa only -> cat[xa * SiLU(xb), xb] -> fc2."""
    t = F.linear(x, sd[prefix + ".fc1.weight"], sd[prefix + ".fc1.bias"])
    c2 = t.shape[-1] // 2
    xa = _to_img(t[..., :c2], h, w)
    xb = t[..., c2:]
    xa = _to_tok(F.conv2d(xa, sd[prefix + ".dw.weight"], sd[prefix + ".dw.bias"], padding=1, groups=c2))
    t = torch.cat([xa * F.silu(xb), xb], dim=-1)
    return F.linear(t, sd[prefix + ".fc2.weight"], sd[prefix + ".fc2.bias"])


def hab(x, sd, prefix, h, w, cfg):
    """HAB.forward, hat_arch.py:217-238 (drop_path is identity in eval); HATX: hatx_arch.py:232-255 (SGFN)."""
    shortcut = x
    xn = _ln(x, sd, prefix + ".norm1")
    conv_x = _to_tok(cab(_to_img(xn, h, w), sd, prefix + ".conv_block"))
    attn_x = _to_tok(conv_attn(_to_img(xn, h, w), sd, prefix + ".esc_attn.core",
                               sd[prefix + ".esc_attn.plk_filter"], cfg["esc_pdim"]))
    x = shortcut + attn_x + conv_x * cfg["conv_scale"]
    ffn = spatial_gate_ffn if cfg.get("arch") == "HATX" else gated_dconv_ffn
    return x + ffn(_ln(x, sd, prefix + ".norm2"), sd, prefix + ".mlp", h, w)


def ocab_bias(sd, prefix, rpi, nq, nk):
    """relative_position_bias gather with wrapping negative indices,
    hat_arch.py:378-382.  Returns (heads, nq, nk)."""
    table = sd[prefix + ".relative_position_bias_table"]
    bias = table[rpi.reshape(-1)].view(nq, nk, -1)
    return bias.permute(2, 0, 1).contiguous()


def ocab(x, sd, prefix, h, w, cfg, heads):
    """OCAB.forward, hat_arch.py:326-393."""
    b, n, c = x.shape
    ws = cfg["window_size"]
    ows = int(ws * cfg["overlap_ratio"]) + ws
    d = c // heads
    scale = cfg.get("qk_scale") or d ** -0.5
    shortcut = x
    xn = _ln(x, sd, prefix + ".norm1")
    y = xn
    if cfg.get("ocab_esc_enable", False):                   # hat_arch.py:336-344
        y = _to_tok(conv_attn(_to_img(xn, h, w), sd, prefix + ".esc_core",
                              sd[prefix + ".esc_plk"], cfg["ocab_esc_pdim"]))
    q = F.linear(xn, sd[prefix + ".q_proj.weight"], sd.get(prefix + ".q_proj.bias"))
    kv = F.linear(y, sd[prefix + ".kv_proj.weight"], sd.get(prefix + ".kv_proj.bias"))
    nwy, nwx = h // ws, w // ws
    # window_partition (hat_arch.py:124-128)
    qw = q.view(b, nwy, ws, nwx, ws, c).permute(0, 1, 3, 2, 4, 5).reshape(b * nwy * nwx, ws * ws, c)
    # nn.Unfold(ows, stride ws, pad (ows-ws)//2) over (B, 2C, H, W), hat_arch.py:359-364; HATX pads by the CEILING
    # (hatx_arch.py:316-322: overlap_ratio 0.6 gives a 25-wide window, 5 before and 4 after)
    pad = (ows - ws + 1) // 2 if cfg.get("arch") == "HATX" else (ows - ws) // 2
    kvu = F.unfold(_to_img(kv, h, w), kernel_size=ows, stride=ws, padding=pad)
    kvu = kvu.view(b, 2, c, ows * ows, nwy * nwx).permute(1, 0, 4, 3, 2).reshape(2, b * nwy * nwx, ows * ows, c)
    kw, vw = kvu[0], kvu[1]
    nb = qw.shape[0]
    qh = qw.view(nb, ws * ws, heads, d).permute(0, 2, 1, 3) * scale
    kh = kw.view(nb, ows * ows, heads, d).permute(0, 2, 1, 3)
    vh = vw.view(nb, ows * ows, heads, d).permute(0, 2, 1, 3)
    attn = qh @ kh.transpose(-2, -1)
    if cfg.get("arch") == "HATX":
        attn = hatx_focus(attn, kw, y, sd, prefix, h, w, cfg, ows, pad)
    attn = attn + ocab_bias(sd, prefix, sd["relative_position_index_OCA"], ws * ws, ows * ows)[None]
    attn = torch.softmax(attn, dim=-1)
    o = (attn @ vh).transpose(1, 2).reshape(nb, ws * ws, c)
    # window_reverse (hat_arch.py:130-134)
    o = o.view(b, nwy, nwx, ws, ws, c).permute(0, 1, 3, 2, 4, 5).reshape(b, h * w, c)
    x = F.linear(o, sd[prefix + ".proj.weight"], sd[prefix + ".proj.bias"]) + shortcut
    m = F.linear(_ln(x, sd, prefix + ".norm2"), sd[prefix + ".mlp.0.weight"], sd[prefix + ".mlp.0.bias"])
    m = F.linear(F.gelu(m), sd[prefix + ".mlp.2.weight"], sd[prefix + ".mlp.2.bias"])
    return x + m


def topk_keep_mask(score, k_keep):
    """Keep the k_keep keys with the largest score; ties at the threshold are broken by the LOWER key index.  (The
    reference calls torch.topk(sorted=False), hatx_arch.py:430, whose choice among equal scores is unspecified; equal
    scores occur only among zero-padded keys, whose score is exactly 0 / tanh(0).)  score: (nb, nk) -> bool (nb, nk)."""
    nk = score.shape[1]
    order = torch.sort(score, dim=1, descending=True, stable=True).indices      # stable: equal scores keep index order
    keep = torch.zeros_like(score, dtype=torch.bool)
    keep.scatter_(1, order[:, :k_keep], True)
    return keep


def hatx_focus(attn, kw, y, sd, prefix, h, w, cfg, ows, pad):
    """OCAB 'focus' options of HATX, hatx_arch.py:405-441: optional per-key focus bias tanh(focus_head(y)) added to the
    logits, optional top-k key pruning (score = focus bias if present, else ||k||_2) with -1e4 on the pruned keys."""
    ws = cfg["window_size"]
    nb, nk = kw.shape[0], kw.shape[1]
    focus = None
    if cfg.get("use_focus_bias", False):
        sal = _conv(F.gelu(_conv(_to_img(y, h, w), sd, prefix + ".focus_head.0", 0)), sd, prefix + ".focus_head.2", 0)
        focus = torch.tanh(F.unfold(sal, kernel_size=ows, stride=ws, padding=pad).transpose(1, 2).reshape(nb, nk))
        attn = attn + focus.view(nb, 1, 1, nk)
    ratio = float(cfg.get("kv_topk_ratio", 1.0))
    if ratio < 1.0:
        k_keep = max(1, int(ratio * nk))
        score = focus if focus is not None else torch.linalg.vector_norm(kw, ord=2, dim=-1)
        keep = topk_keep_mask(score, k_keep)
        attn = attn.masked_fill(~keep.view(nb, 1, 1, nk), -1e4)
    return attn


def rhag(x, sd, prefix, h, w, cfg, depth, heads):
    """RHAG.forward (hat_arch.py:555-556) over AttenBlocks.forward (:466-482)."""
    y = x
    for i in range(depth):
        y = hab(y, sd, f"{prefix}.residual_group.blocks.{i}", h, w, cfg)
    y = ocab(y, sd, f"{prefix}.residual_group.overlap_attn", h, w, cfg, heads)
    if cfg.get("resi_connection", "1conv") == "1conv":
        y = _to_tok(_conv(_to_img(y, h, w), sd, prefix + ".conv", 1))
    return y + x


def upsample(x, sd, scale):
    """Upsample, hat_arch.py:593-605."""
    if scale & (scale - 1) == 0:
        for i in range(int(math.log2(scale))):
            x = F.pixel_shuffle(_conv(x, sd, f"upsample.{2 * i}", 1), 2)
    elif scale == 3:
        x = F.pixel_shuffle(_conv(x, sd, "upsample.0", 1), 3)
    else:
        raise ValueError(f"scale {scale} is not supported. Supported scales: 2^n and 3.")
    return x


def forward_features(x, sd, cfg):
    """HAT.forward_features, hat_arch.py:828-846 (mask skipped, see header)."""
    b, c, h, w = x.shape
    ws = cfg["window_size"]
    if h % ws or w % ws or h < ws or w < ws:
        raise RuntimeError(f"input {h}x{w} is not a multiple of window_size {ws}")
    t = _to_tok(x)
    if cfg.get("patch_norm", True):
        t = _ln(t, sd, "patch_embed.norm")
    if cfg.get("ape", False):
        t = t + sd["absolute_pos_embed"]
    for i, depth in enumerate(cfg["depths"]):
        t = rhag(t, sd, f"layers.{i}", h, w, cfg, depth, cfg["num_heads"][i])
    t = _ln(t, sd, "norm")
    return _to_img(t, h, w)


def hat_forward(x, sd, cfg):
    """HAT.forward, hat_arch.py:848-859.  x: (B, in_chans, H, W) float32 CPU."""
    cfg = default_cfg(**cfg)
    if cfg["in_chans"] == 3:
        mean = torch.tensor(RGB_MEAN, dtype=x.dtype).view(1, 3, 1, 1)
    else:
        mean = torch.zeros(1, 1, 1, 1, dtype=x.dtype)
    rng = cfg["img_range"]
    with torch.no_grad():
        x = (x - mean) * rng
        if cfg["upsampler"] == "pixelshuffle":
            x = _conv(x, sd, "conv_first", 1)
            f = forward_features(x, sd, cfg)
            if cfg.get("resi_connection", "1conv") == "1conv":
                f = _conv(f, sd, "conv_after_body", 1)
            x = f + x
            x = F.leaky_relu(_conv(x, sd, "conv_before_upsample.0", 1), 0.01)
            x = _conv(upsample(x, sd, cfg["upscale"]), sd, "conv_last", 1)
        return x / rng + mean


def tiled_forward(img, sd, cfg, tile_size, tile_pad, forward=None):
    """HATModel.pre_process -> tile_process -> post_process
    (hat_model.py:16-26, :40-112).  img: (B,3,H,W)."""
    cfg = default_cfg(**cfg)
    scale = cfg["upscale"]
    forward = forward or (lambda t: hat_forward(t, sd, cfg))
    _, _, h0, w0 = img.shape
    ph, pw = mod_pad(h0, w0, cfg["window_size"])
    img = F.pad(img, (0, pw, 0, ph), "reflect")
    b, c, h, w = img.shape
    out = img.new_zeros(b, c, h * scale, w * scale)
    for r in tile_rects(h, w, tile_size, tile_pad, scale):
        o = forward(img[:, :, r["in_y0"]:r["in_y1"], r["in_x0"]:r["in_x1"]])
        out[:, :, r["out_y0"]:r["out_y1"], r["out_x0"]:r["out_x1"]] = \
            o[:, :, r["crop_y0"]:r["crop_y1"], r["crop_x0"]:r["crop_x1"]]
    return out[:, :, :h * scale - ph * scale, :w * scale - pw * scale]


# --------------------------------------------------------------------------
# deterministic state_dict generator (independent of the reference's init)
# --------------------------------------------------------------------------
def randomized_state_dict(template: dict, seed: int):
    """Redraw every floating parameter of `template` (a state_dict giving key
    order/shapes): weights N(0, 0.05), biases N(0, 0.02), LayerNorm weight
    1+N(0, 0.1), bias table N(0, 0.2).  Integer buffers are kept.  Random-init
    HAT has zero dynamic kernels and identity LN, so parity also needs this
    (SURVEY.md 0.5)."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k, v in template.items():
        if not v.is_floating_point():
            out[k] = v.clone()
            continue
        r = torch.randn(v.shape, generator=g, dtype=torch.float32)
        if "norm" in k.split(".")[-2] and k.endswith(".weight"):
            out[k] = 1.0 + 0.1 * r
        elif k.endswith("relative_position_bias_table"):
            out[k] = 0.2 * r
        elif k.endswith(".bias"):
            out[k] = 0.02 * r
        elif k.endswith("plk_filter") or k.endswith("esc_plk"):
            out[k] = 0.02 * r
        elif "dwc_proj" in k:
            out[k] = 0.2 * r
        else:
            out[k] = 0.05 * r
    return out


def scale_tail(sd: dict, tail_scale: float):
    """`conv_last.{weight,bias}` x float32(tail_scale): y' = tail_scale * (y - mean) + mean, so a randomised
    state_dict gives image-range outputs (the north_star's bars are stated at peak 1.0)."""
    out = dict(sd)
    s = torch.tensor(float(tail_scale), dtype=torch.float32)
    for k in ("conv_last.weight", "conv_last.bias"):
        out[k] = sd[k] * s
    return out


def fresh_state_dict(hat_cls, cfg: dict, seed: int):
    """The reference-style initialisation (hat_arch.py:761-768 + PyTorch conv defaults) of `hat_cls(**cfg)`
    under torch.manual_seed(seed), as a detached CPU state_dict."""
    torch.manual_seed(seed)
    net = hat_cls(**cfg)
    return {k: v.detach().clone() for k, v in net.state_dict().items()}


def state_dict_checksum(sd: dict) -> float:
    """Order-dependent float64 checksum of the floating entries (pins seeded weights across machines)."""
    tot = 0.0
    for i, (k, v) in enumerate(sd.items()):
        if v.is_floating_point():
            tot += (i % 7 + 1) * v.double().abs().sum().item()
    return tot
