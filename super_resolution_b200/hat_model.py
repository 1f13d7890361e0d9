"""Tile-mode inference around the HAT forward: the caller side of the hot path.

Mirrors `HATModel.pre_process / process / tile_process / post_process` of the reference
(HAT/hat/models/hat_model.py:16-112) with the same names, option keys and index arithmetic, but
  * the geometry comes from the C ABI (`hatb200_tile_plan`), bit-exact with hat_model.py:52-108;
  * reflect padding, crop-and-place run as CUDA kernels of libhatb200 (`hatb200_reflect_pad`,
    `hatb200_crop_place`) on the current stream -- no host round trip, no per-tile print;
  * tiles of equal padded shape are stacked along the batch axis into ONE forward (every per-sample
    statistic of the network -- ECA pool, dynamic-kernel pool -- is confined to a sample, SURVEY.md 0.2, so a
    stacked forward equals the reference's one-tile-at-a-time loop);
  * with a process group, tiles are sharded over the ranks (longest-processing-time-first on padded area,
    SURVEY.md 8e) and the only collective is the gather of the cropped HR cores to `dst`
    (send/recv over NCCL/NVLink; gloo in the CPU tests).

PyTorch is plumbing here (device buffers, streams, torch.distributed); all arithmetic happens in libhatb200.
There is no CPU path: the image must be a CUDA tensor.
"""
import ctypes as C
from collections import OrderedDict

import torch

from . import _lib


# --------------------------------------------------------------------------------------------------
# host-side planning (pure integer bookkeeping; runs anywhere)
# --------------------------------------------------------------------------------------------------
def mod_pad(h: int, w: int, window_size: int):
    """hat_model.py:18-25: rows / columns to add so that both are multiples of window_size."""
    ph = (window_size - h % window_size) % window_size
    pw = (window_size - w % window_size) % window_size
    return ph, pw


# One stacked forward costs about as much as this many LR pixels of tile area on top of the area it processes
# (B200, HAT-L tiles of 288-320 px, tools/time_tiles.py: T = 0.27 us/px * area + ~5.5 ms per forward -- launch ramps of
# ~670 kernels and the last-wave quantisation of 650-800 work tiles on 148 SMs; profiles/r02_tiled_scaling.md)
FORWARD_COST_PX = 20000


def shard_cost(rects, idxs, forward_cost_px=FORWARD_COST_PX):
    """Cost model of one rank's share: padded input area + a fixed cost per stacked forward (= per distinct shape)."""
    shapes = {(rects[i]["in_y1"] - rects[i]["in_y0"], rects[i]["in_x1"] - rects[i]["in_x0"]) for i in idxs}
    return sum(h * w for h, w in ((rects[i]["in_y1"] - rects[i]["in_y0"], rects[i]["in_x1"] - rects[i]["in_x0"])
                                  for i in idxs)) + forward_cost_px * len(shapes)


_SHARD_CACHE = {}


def shard_tiles(rects, world_size: int, forward_cost_px: int = FORWARD_COST_PX):
    """Memoised `_shard_tiles` (the local search costs tens of milliseconds of host time for 16 tiles on 8 ranks;
    an image stream of one size plans once)."""
    key = (tuple((r["in_y1"] - r["in_y0"], r["in_x1"] - r["in_x0"]) for r in rects), world_size, forward_cost_px)
    if key not in _SHARD_CACHE:
        if len(_SHARD_CACHE) > 64:
            _SHARD_CACHE.clear()
        _SHARD_CACHE[key] = _shard_tiles(rects, world_size, forward_cost_px)
    return [list(o) for o in _SHARD_CACHE[key]]


def _shard_tiles(rects, world_size: int, forward_cost_px: int = FORWARD_COST_PX):
    """Assign tiles to ranks: longest-processing-time-first on the padded input area, then a deterministic local
    search (single-tile moves and pairwise swaps that lower the maximum, then the sum of squares, of `shard_cost`) --
    same-shape tiles share one stacked forward, so a rank with two tiles of one shape beats one with two shapes.

    Deterministic (ties broken by the reference tile order hat_model.py:76), so every rank computes the same
    assignment without communicating.  Returns a list of `world_size` lists of tile indices, each in ascending
    (reference) order."""
    area = lambda r: (r["in_y1"] - r["in_y0"]) * (r["in_x1"] - r["in_x0"])
    shape = lambda r: (r["in_y1"] - r["in_y0"], r["in_x1"] - r["in_x0"])
    if world_size > 1 and len(rects) > 32 and forward_cost_px > 0:
        # Many units per rank (a batch of images): shape-major contiguous partition.  Units sorted by (shape, index) are
        # cut into world_size runs of equal area, so a rank holds one or two shapes and its stacks are large (B200:
        # 0.20 us per LR pixel in stacks of ten 320x320 tiles against 0.35 when every rank gets all nine shapes of a
        # 540x960 image in forwards of one or two tiles; profiles/r02_tiled_scaling.md).
        order = sorted(range(len(rects)), key=lambda i: (-area(rects[i]), shape(rects[i]), i))
        total = sum(area(r) for r in rects)
        owner = [[] for _ in range(world_size)]
        acc, k = 0, 0
        for i in order:
            a = area(rects[i])
            # move on when this unit's midpoint lies beyond the current rank's share of the total area
            while k < world_size - 1 and acc + a / 2 > total * (k + 1) / world_size:
                k += 1
            owner[k].append(i)
            acc += a
        return [sorted(o) for o in owner]
    order = sorted(range(len(rects)), key=lambda i: (-area(rects[i]), i))
    load = [0] * world_size
    owner = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        owner[r].append(i)
        load[r] += area(rects[i])
    if world_size == 1 or forward_cost_px <= 0:
        return [sorted(o) for o in owner]
    cost = lambda o: shard_cost(rects, o, forward_cost_px)
    key = lambda ow: (max(cost(o) for o in ow), sum(cost(o) ** 2 for o in ow))
    best = key(owner)
    for _ in range(4 * len(rects)):                 # each accepted step strictly lowers `key`: terminates
        cand, cand_key = None, best
        for a in range(world_size):
            for b in range(world_size):
                if a == b:
                    continue
                for i in owner[a]:
                    trial = [list(o) for o in owner]            # move i: a -> b
                    trial[a].remove(i); trial[b].append(i)
                    k = key(trial)
                    if k < cand_key:
                        cand, cand_key = trial, k
                    if a < b:
                        for j in owner[b]:                      # swap i <-> j
                            trial = [list(o) for o in owner]
                            trial[a].remove(i); trial[b].remove(j); trial[a].append(j); trial[b].append(i)
                            k = key(trial)
                            if k < cand_key:
                                cand, cand_key = trial, k
        if cand is None:
            break
        owner, best = cand, cand_key
    return [sorted(o) for o in owner]


def group_by_shape(rects, indices):
    """Same-shape tiles -> one stacked forward.  Ordered dict {(h, w): [tile indices]}."""
    groups = OrderedDict()
    for i in indices:
        r = rects[i]
        groups.setdefault((r["in_y1"] - r["in_y0"], r["in_x1"] - r["in_x0"]), []).append(i)
    return groups


# hatb200_forward indexes tokens with 32 bits (csrc/forward.cu check_geometry: B*H*W*16*64 < 2^31); a stacked forward
# also needs ~1 KB of workspace per token, so stacks are capped well below that limit by default (B200, HAT x4,
# tools/time_scaling.py: 0.557 / 0.548 / 0.547 ms per 64x64 patch at 262k / 524k / 1M tokens -- no cliff up to the cap)
MAX_FORWARD_TOKENS = (1 << 21) - 1
DEFAULT_STACK_TOKENS = 1 << 20


def split_stacks(idxs, th, tw, batch, max_tokens=DEFAULT_STACK_TOKENS):
    """Cut one shape group into stacks of at most `max_tokens` LR pixels per forward (at least one tile each; the
    reference runs one tile at a time and therefore handles any image size, hat_model.py:56-88)."""
    per = max(1, min(max_tokens, MAX_FORWARD_TOKENS) // max(1, th * tw * batch))
    return [idxs[k:k + per] for k in range(0, len(idxs), per)]


def core_shape(rect, batch, channels):
    return (batch, channels, rect["out_y1"] - rect["out_y0"], rect["out_x1"] - rect["out_x0"])


# --------------------------------------------------------------------------------------------------
# the one collective: cropped HR cores -> dst
# --------------------------------------------------------------------------------------------------
def gather_cores(cores, rects, owners, batch, channels, dtype, device, dst=0, group=None):
    """cores: {tile index: (B, C, rh, rw) tensor} owned by this rank.  Returns the full
    {tile index: tensor} on `dst` (None elsewhere).  One flat message per (rank -> dst) pair; the message
    layout is implied by the shared plan, so no sizes are exchanged."""
    import torch.distributed as dist
    rank = dist.get_rank(group)
    world = dist.get_world_size(group)
    if world == 1:
        return dict(cores)
    flat = None
    if rank != dst and owners[rank]:
        flat = torch.cat([cores[i].reshape(-1) for i in owners[rank]])
    bufs = exchange_flat(flat, rects, owners, batch, channels, dtype, device, dst, group)
    if rank != dst:
        return None
    out = dict(cores)
    for r, buf in bufs.items():
        for i, view in flat_views(buf, rects, owners[r], batch, channels).items():
            out[i] = view
    return out


def flat_numel(rects, idxs, batch, channels):
    return sum(int(torch.Size(core_shape(rects[i], batch, channels)).numel()) for i in idxs)


def flat_views(buf, rects, idxs, batch, channels):
    """{tile index: (B, C, rh, rw) view} of one rank's flat message (tiles in ascending order, each contiguous)."""
    out, off = {}, 0
    for i in idxs:
        shp = core_shape(rects[i], batch, channels)
        n = int(torch.Size(shp).numel())
        out[i] = buf[off:off + n].view(shp)
        off += n
    return out


def exchange_flat(flat, rects, owners, batch, channels, dtype, device, dst=0, group=None):
    """The one collective of tile mode: every rank other than `dst` sends its flat message of cropped HR cores,
    `dst` receives one buffer per peer (returned as {rank: buffer}; {} elsewhere).  NCCL send / recv over NVLink."""
    import torch.distributed as dist
    rank = dist.get_rank(group)
    world = dist.get_world_size(group)
    peer = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)
    ops, bufs = [], {}
    if rank == dst:
        for r in range(world):
            if r == dst or not owners[r]:
                continue
            bufs[r] = torch.empty(flat_numel(rects, owners[r], batch, channels), dtype=dtype, device=device)
            ops.append(dist.P2POp(dist.irecv, bufs[r], peer(r), group))
    elif flat is not None:
        ops.append(dist.P2POp(dist.isend, flat, peer(dst), group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return bufs


# --------------------------------------------------------------------------------------------------
# device-side pipeline
# --------------------------------------------------------------------------------------------------
def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream(t):
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _require_cuda(t):
    if not t.is_cuda:
        raise RuntimeError("super_resolution_b200.HATModel has no CPU path: the image must be a CUDA tensor")


def reflect_pad(img, ph, pw):
    """F.pad(img, (0, pw, 0, ph), 'reflect') (hat_model.py:26) as one kernel."""
    _require_cuda(img)
    if ph == 0 and pw == 0:
        return img
    B, Cn, h, w = img.shape
    if ph >= h or pw >= w:
        raise RuntimeError("Padding size should be less than the corresponding input dimension")   # torch's wording
    img = img.contiguous()
    out = torch.empty((B, Cn, h + ph, w + pw), dtype=img.dtype, device=img.device)
    _lib.check(_lib.lib().hatb200_reflect_pad(_ptr(img), _lib.dtype_code(img.dtype), B, Cn, h, w, _ptr(out), ph, pw,
                                              _stream(img)))
    return out


def crop_place(src, cy, cx, dst, oy, ox, rh, rw):
    """dst[:, :, oy:oy+rh, ox:ox+rw] = src[:, :, cy:cy+rh, cx:cx+rw] (hat_model.py:106-108) as one kernel."""
    B, Cn, sh, sw = src.shape
    assert src.is_contiguous() and dst.is_contiguous() and dst.shape[:2] == src.shape[:2]
    _lib.check(_lib.lib().hatb200_crop_place(_ptr(src), _lib.dtype_code(src.dtype), B, Cn, sh, sw, cy, cx, _ptr(dst),
                                             dst.shape[2], dst.shape[3], oy, ox, rh, rw, _stream(src)))


def tensor2img(tensor, rgb2bgr=True, out_type=None, min_max=(0, 1), to_numpy=True):
    """Device-side `basicsr.utils.tensor2img` (basicsr/utils/img_util.py:38-95): clamp, rescale, x255, round, uint8,
    CHW -> HWC, RGB -> BGR in ONE kernel (`hatb200_tensor2img`) before the device-to-host copy -- the reference copies the
    fp32 image to the host first and does this in numpy / cv2.

    tensor: CUDA tensor (1,C,H,W), (C,H,W) or (H,W), or a list of them (a list returns a list, as the reference).
    out_type: None / numpy.uint8 only (the float output of the reference is its input rescaled; no kernel needed).
    Returns a numpy (H,W,C) / (H,W) uint8 array, or the device uint8 tensor when to_numpy=False."""
    import numpy as np
    if out_type is not None and out_type is not np.uint8:
        raise NotImplementedError("tensor2img on device produces uint8 only")
    if not (torch.is_tensor(tensor) or (isinstance(tensor, (list, tuple)) and all(torch.is_tensor(t) for t in tensor))):
        raise TypeError(f'tensor or list of tensors expected, got {type(tensor)}')      # the reference's wording
    many = not torch.is_tensor(tensor)
    result = []
    for t in (tensor if many else [tensor]):
        if not t.is_cuda:
            raise RuntimeError("super_resolution_b200.tensor2img: CUDA tensor required (there is no CPU path)")
        t = t.squeeze(0)
        if t.dim() == 4:
            raise NotImplementedError("tensor2img: 4-D mini-batches go through torchvision make_grid in the reference")
        if t.dim() not in (2, 3):
            raise TypeError(f'Only support 4D, 3D or 2D tensor. But received with dimension: {t.dim()}')
        if t.dtype not in (torch.float32, torch.bfloat16):
            t = t.float()
        t = t.contiguous()
        Cn, H, W = (1, *t.shape) if t.dim() == 2 else t.shape
        out = torch.empty((H, W, Cn), dtype=torch.uint8, device=t.device)
        _lib.check(_lib.lib().hatb200_tensor2img(_ptr(t), _lib.dtype_code(t.dtype), 1, Cn, H, W, float(min_max[0]),
                                                 float(min_max[1]), int(bool(rgb2bgr)), _ptr(out), _stream(t)))
        if Cn == 1:
            out = out[:, :, 0]                      # gray image: (H, W), as np.squeeze(axis=2) in the reference
        result.append(out.cpu().numpy() if to_numpy else out)
    return result if many else result[0]


def calculate_psnr(img, img2, crop_border, input_order='HWC', test_y_channel=False):
    """Device-side `basicsr.metrics.calculate_psnr` (basicsr/metrics/psnr_ssim.py:12-50) for two uint8 CUDA images
    with range [0, 255] (HWC, BGR when 3 channels -- what `tensor2img` returns).  One reduction kernel + a fixed-order
    finish; only two doubles cross to the host."""
    import math
    assert img.shape == img2.shape, (f'Image shapes are different: {img.shape}, {img2.shape}.')
    if input_order not in ['HWC', 'CHW']:
        raise ValueError(f'Wrong input_order {input_order}. Supported input_orders are "HWC" and "CHW"')
    if not (img.is_cuda and img2.is_cuda and img.dtype == torch.uint8 and img2.dtype == torch.uint8):
        raise RuntimeError("super_resolution_b200.calculate_psnr: uint8 CUDA tensors required")
    if img.dim() == 2:
        img, img2 = img[..., None], img2[..., None]
    elif input_order == 'CHW':
        img, img2 = img.permute(1, 2, 0), img2.permute(1, 2, 0)
    img, img2 = img.contiguous(), img2.contiguous()
    H, W, Cn = img.shape
    ws = torch.empty(1024, dtype=torch.float64, device=img.device)
    out = torch.empty(2, dtype=torch.float64, device=img.device)
    _lib.check(_lib.lib().hatb200_psnr_u8(_ptr(img), _ptr(img2), H, W, Cn, int(crop_border), int(bool(test_y_channel)),
                                          _ptr(ws), _ptr(out), _stream(img)))
    sse, count = out.cpu().tolist()
    if sse == 0:
        return float('inf')
    return 10. * math.log10(255. * 255. / (sse / count))


def calculate_ssim(img, img2, crop_border, input_order='HWC', test_y_channel=False):
    """Device-side `basicsr.metrics.calculate_ssim` (basicsr/metrics/psnr_ssim.py:85-129, `_ssim` :170-198) for two uint8
    CUDA images with range [0, 255] (HWC, BGR when 3 channels): float64, 11x11 Gaussian window over the valid region,
    mean over channels.  One stencil kernel + a fixed-order finish; two doubles cross to the host."""
    assert img.shape == img2.shape, (f'Image shapes are different: {img.shape}, {img2.shape}.')
    if input_order not in ['HWC', 'CHW']:
        raise ValueError(f'Wrong input_order {input_order}. Supported input_orders are "HWC" and "CHW"')
    if not (img.is_cuda and img2.is_cuda and img.dtype == torch.uint8 and img2.dtype == torch.uint8):
        raise RuntimeError("super_resolution_b200.calculate_ssim: uint8 CUDA tensors required")
    if img.dim() == 2:
        img, img2 = img[..., None], img2[..., None]
    elif input_order == 'CHW':
        img, img2 = img.permute(1, 2, 0), img2.permute(1, 2, 0)
    img, img2 = img.contiguous(), img2.contiguous()
    H, W, Cn = img.shape
    ws = torch.empty(4096, dtype=torch.float64, device=img.device)
    out = torch.empty(2, dtype=torch.float64, device=img.device)
    _lib.check(_lib.lib().hatb200_ssim_u8(_ptr(img), _ptr(img2), H, W, Cn, int(crop_border), int(bool(test_y_channel)),
                                          _ptr(ws), _ptr(out), _stream(img)))
    total, count = out.cpu().tolist()
    return total / count


def save_network(net, save_path, param_key='params'):
    """Checkpoint file in the reference's on-disk format (`BaseModel.save_network`, basicsr/models/base_model.py:207-252):
    `{param_key: state_dict}` with CPU tensors, `module.` prefixes (DataParallel / DDP wrappers) stripped.  `net` and
    `param_key` may be parallel lists, as in the reference (`['params', 'params_ema']`)."""
    nets = net if isinstance(net, list) else [net]
    keys = param_key if isinstance(param_key, list) else [param_key]
    assert len(nets) == len(keys), 'The lengths of net and param_key should be the same.'
    save_dict = {}
    for net_, key_ in zip(nets, keys):
        net_ = getattr(net_, 'module', net_)              # get_bare_model
        save_dict[key_] = {(k[7:] if k.startswith('module.') else k): v.cpu() for k, v in net_.state_dict().items()}
    torch.save(save_dict, save_path)


def load_network(net, load_path, strict=True, param_key='params'):
    """`BaseModel.load_network` (basicsr/models/base_model.py:289-315): reads a `.pth` written by the reference's
    `save_network`, falls back from a missing `params_ema` to `params`, strips `module.` prefixes and loads with
    `load_state_dict` -- so checkpoints trained with the reference drop into the B200 `HAT` unchanged.  With
    strict=False, keys whose sizes differ are skipped (base_model.py:281-287)."""
    net = getattr(net, 'module', net)
    load_net = torch.load(load_path, map_location=lambda storage, loc: storage, weights_only=True)
    if param_key is not None:
        if param_key not in load_net and 'params' in load_net:
            param_key = 'params'
        load_net = load_net[param_key]
    load_net = {(k[7:] if k.startswith('module.') else k): v for k, v in load_net.items()}
    if not strict:
        crt = net.state_dict()
        for k in list(load_net.keys()):
            if k in crt and crt[k].size() != load_net[k].size():
                load_net[k + '.ignore'] = load_net.pop(k)
    return net.load_state_dict(load_net, strict=strict)


class HATModel:
    """Inference half of the reference `HATModel` (hat_model.py:13-185) around a B200-native `HAT`.

    Two ways to build it:
      * `HATModel(opt)` -- the reference's signature, what `basicsr.models.build_model(opt)` calls
        (`MODEL_REGISTRY.get(opt['model_type'])(opt)`, basicsr/models/__init__.py:17-28): builds the network from
        `opt['network_g']` (type 'HAT'), moves it to the GPU and loads `opt['path']['pretrain_network_g']` with
        `param_key_g` / `strict_load_g` (sr_model.py:18-32).  `hat/test.py` therefore runs unchanged once this
        package is imported in place of `hat.models`.  Forward-only: `opt['is_train']` raises.
      * `HATModel(net, opt)` -- around a network that already exists.

    opt: the reference's option dict; the keys read are the ones the reference reads:
    `opt['network_g']['window_size']`, `opt['scale']`, and optionally `opt['tile']['tile_size']`,
    `opt['tile']['tile_pad']` (presence of `opt['tile']` selects tile mode, hat_model.py:139-141); for
    `nondist_validation` also `opt['val']`, `opt['path']['visualization']`, `opt['name']`, `opt['is_train']`.
    group/dst: optional torch.distributed process group for tile sharding and the rank that receives the image.

    Sharding is opt-in, as the reference validates on one rank (sr_model.py:180-182 dist_validation ->
    nondist_validation on rank 0): tiles are spread over ranks only when a process group is passed or
    `opt['tile']['shard']` is true -- then EVERY rank of the group must call `tile_process` with the SAME image.
    With an initialised default group and neither of those, each caller computes all of its tiles itself.
    `opt['tile']['max_stack_tokens']` (default 2^20 LR pixels) caps how many same-shape tiles share one forward.
    """

    def __init__(self, net_or_opt, opt=None, group=None, dst=0):
        if opt is None and isinstance(net_or_opt, dict):
            opt = net_or_opt
            net = self._build_from_opt(opt)
        else:
            net = net_or_opt
        self.net_g = net
        self.opt = opt
        self.group = group
        self.dst = dst
        self.is_train = bool(opt.get('is_train', False))
        self.scale = opt.get('scale', 1)
        self.mod_pad_h = self.mod_pad_w = 0
        self.lq = self.img = self.output = None
        self.record_events = False      # bench: CUDA event after this rank's last forward / crop, before the gather
        self.events = {}
        try:
            self.device = next(net.parameters()).device
        except (StopIteration, AttributeError):
            self.device = torch.device('cuda')

    @staticmethod
    def _build_from_opt(opt):
        """SRModel.__init__ (sr_model.py:18-32): build_network -> model_to_device -> load_network."""
        import copy
        if opt.get('is_train', False):
            raise NotImplementedError("super_resolution_b200.HATModel is inference-only (is_train must be false)")
        if opt.get('num_gpu', 1) == 0 or not torch.cuda.is_available():
            raise RuntimeError("super_resolution_b200.HATModel has no CPU path: a CUDA device is required")
        from .hat_arch import HAT
        net_opt = copy.deepcopy(opt['network_g'])
        typ = net_opt.pop('type', 'HAT')
        if typ not in ('HAT', 'HAT_b200'):
            raise NotImplementedError(f"network_g type '{typ}': this package implements the reference's 'HAT'")
        net = HAT(**net_opt).eval().to(torch.device('cuda', torch.cuda.current_device()))
        path = opt.get('path', {}) or {}
        load_path = path.get('pretrain_network_g', None)
        if load_path is not None:
            load_network(net, load_path, path.get('strict_load_g', True), path.get('param_key_g', 'params'))
        return net

    # -- reference API -------------------------------------------------------------------------
    def feed_data(self, data):                      # sr_model.py:87-90
        if isinstance(data, dict):
            self.lq = data['lq'].to(self.device)
            if 'gt' in data:
                self.gt = data['gt'].to(self.device)
        else:
            self.lq = data

    def pre_process(self):                          # hat_model.py:16-26
        window_size = self.opt['network_g']['window_size']
        self.scale = self.opt.get('scale', 1)
        _, _, h, w = self.lq.size()
        self.mod_pad_h, self.mod_pad_w = mod_pad(h, w, window_size)
        self.img = reflect_pad(self.lq, self.mod_pad_h, self.mod_pad_w)

    def process(self):                              # hat_model.py:28-38
        self.net_g.eval()
        with torch.no_grad():
            self.output = self.net_g(self.img)

    def tile_process(self):                         # hat_model.py:40-108
        img = self.img
        _require_cuda(img)
        batch, channel, height, width = img.shape
        ts, tp = self.opt['tile']['tile_size'], self.opt['tile']['tile_pad']
        rects = _lib.tile_plan(height, width, ts, tp, self.scale)
        world, rank = 1, 0
        if self.group is not None or (self.opt['tile'].get('shard') and _dist_ready()):
            import torch.distributed as dist
            world, rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        # The unit of work is one tile of ONE sample (every statistic of the network is per sample):
        # unit u = tile * batch + sample, in the reference's loop order (tile outside, hat_model.py:56-76).  A batch of
        # images therefore shards over the ranks together with its tiles, and stacks never exceed the token budget.
        units = [dict(rects[i], tile=i, sample=b) for i in range(len(rects)) for b in range(batch)]
        owners = shard_tiles(units, world)
        max_tokens = int(self.opt['tile'].get('max_stack_tokens', DEFAULT_STACK_TOKENS))
        self.net_g.eval()
        is_dst = world == 1 or rank == self.dst
        # "start with black image" (hat_model.py:50): every output pixel is covered by exactly one core.  On the
        # consumer rank the cores of its own units go straight from the network output into the image; the other
        # ranks crop theirs straight into ONE flat send buffer (the message layout is implied by the shared plan).
        if is_dst:
            self.output = torch.empty((batch, channel, height * self.scale, width * self.scale), dtype=img.dtype,
                                      device=img.device)
            send, views = None, {}
        else:
            self.output = None
            send = torch.empty(flat_numel(units, owners[rank], 1, channel), dtype=img.dtype, device=img.device)
            views = flat_views(send, units, owners[rank], 1, channel)
        # the stacked forward's result is cropped right away on the same stream: take the engine's own output buffer
        # instead of a fresh copy of it
        eng = self.net_g.engine() if hasattr(self.net_g, 'engine') else None
        view0 = eng.output_view if eng is not None else False
        if eng is not None:
            eng.output_view = True
        try:
            self._run_stacks(img, units, owners[rank], channel, max_tokens, is_dst, views)
        finally:
            if eng is not None:
                eng.output_view = view0
        if self.record_events:
            self.events['compute_end'] = torch.cuda.Event(enable_timing=True)
            self.events['compute_end'].record()
        if world > 1:
            bufs = exchange_flat(send, units, owners, 1, channel, img.dtype, img.device, self.dst, self.group)
            for r, buf in bufs.items():
                for u, core in flat_views(buf, units, owners[r], 1, channel).items():
                    q = units[u]
                    crop_place(core, 0, 0, self.output[q["sample"]:q["sample"] + 1], q["out_y0"], q["out_x0"],
                               core.shape[2], core.shape[3])

    def _run_stacks(self, img, units, mine, channel, max_tokens, is_dst, views):
        with torch.no_grad():
            stacks = [((th, tw), part) for (th, tw), idxs in group_by_shape(units, mine).items()
                      for part in split_stacks(idxs, th, tw, 1, max_tokens)]
            for (th, tw), idxs in stacks:
                stack = torch.empty((len(idxs), channel, th, tw), dtype=img.dtype, device=img.device)
                for k, u in enumerate(idxs):        # extract (hat_model.py:77): one strided copy kernel per unit
                    r = units[u]
                    crop_place(img[r["sample"]:r["sample"] + 1], r["in_y0"], r["in_x0"], stack[k:k + 1], 0, 0, th, tw)
                out = self.net_g(stack)
                for k, u in enumerate(idxs):        # crop the padding off and place (hat_model.py:106-108)
                    r = units[u]
                    rh, rw = r["out_y1"] - r["out_y0"], r["out_x1"] - r["out_x0"]
                    if is_dst:
                        crop_place(out[k:k + 1], r["crop_y0"], r["crop_x0"], self.output[r["sample"]:r["sample"] + 1],
                                   r["out_y0"], r["out_x0"], rh, rw)
                    else:
                        crop_place(out[k:k + 1], r["crop_y0"], r["crop_x0"], views[u], 0, 0, rh, rw)

    def post_process(self):                         # hat_model.py:110-112
        if self.output is None:
            return
        _, _, h, w = self.output.size()
        rh, rw = h - self.mod_pad_h * self.scale, w - self.mod_pad_w * self.scale
        if rh == h and rw == w:
            return
        out = torch.empty(self.output.shape[:2] + (rh, rw), dtype=self.output.dtype, device=self.output.device)
        crop_place(self.output, 0, 0, out, 0, 0, rh, rw)
        self.output = out

    def get_current_visuals(self):                  # sr_model.py:266-272
        out_dict = OrderedDict()
        out_dict['lq'] = self.lq.detach().cpu()
        out_dict['result'] = self.output.detach().cpu()
        if hasattr(self, 'gt'):
            out_dict['gt'] = self.gt.detach().cpu()
        return out_dict

    def dist_validation(self, dataloader, current_iter, tb_logger, save_img):      # sr_model.py:180-182
        if self.opt.get('rank', 0) == 0:
            self.nondist_validation(dataloader, current_iter, tb_logger, save_img)

    def validation(self, dataloader, current_iter, tb_logger, save_img=False):     # base_model.py:41-53
        if self.opt.get('dist', False):
            self.dist_validation(dataloader, current_iter, tb_logger, save_img)
        else:
            self.nondist_validation(dataloader, current_iter, tb_logger, save_img)

    def nondist_validation(self, dataloader, current_iter, tb_logger, save_img):
        """hat_model.py:114-185, image by image: feed -> pre_process -> tile_process / process -> post_process ->
        tensor2img -> metrics -> imwrite.  The HR image never leaves the GPU as floats: `tensor2img` and the PSNR /
        SSIM metrics (`calculate_psnr`, `calculate_ssim`) run as kernels of libhatb200 on it; only the uint8 HWC image
        (for imwrite) and two doubles per metric cross to the host.  Other metric types go to basicsr's registry."""
        from os import path as osp
        dataset_name = dataloader.dataset.opt['name']
        with_metrics = self.opt['val'].get('metrics') is not None
        if with_metrics:
            if not hasattr(self, 'metric_results'):
                self.metric_results = {metric: 0 for metric in self.opt['val']['metrics'].keys()}
            self._initialize_best_metric_results(dataset_name)
            self.metric_results = {metric: 0 for metric in self.metric_results}
        idx = -1
        for idx, val_data in enumerate(dataloader):
            img_name = osp.splitext(osp.basename(val_data['lq_path'][0]))[0]
            self.feed_data(val_data)
            self.pre_process()
            if 'tile' in self.opt:
                self.tile_process()
            else:
                self.process()
            self.post_process()
            sr_dev = tensor2img([self.output], to_numpy=False)[0]
            gt_dev = None
            if hasattr(self, 'gt'):
                gt_dev = tensor2img([self.gt], to_numpy=False)[0]
                del self.gt
            del self.lq
            self.output = None
            if save_img:
                if self.opt.get('is_train', False):
                    save_img_path = osp.join(self.opt['path']['visualization'], img_name, f'{img_name}_{current_iter}.png')
                elif self.opt['val'].get('suffix'):
                    save_img_path = osp.join(self.opt['path']['visualization'], dataset_name,
                                             f'{img_name}_{self.opt["val"]["suffix"]}.png')
                else:
                    save_img_path = osp.join(self.opt['path']['visualization'], dataset_name,
                                             f'{img_name}_{self.opt["name"]}.png')
                _imwrite(sr_dev.cpu().numpy(), save_img_path)
            if with_metrics:
                for name, opt_ in self.opt['val']['metrics'].items():
                    self.metric_results[name] += _device_metric(sr_dev, gt_dev, opt_)
        if with_metrics and idx >= 0:
            for metric in self.metric_results.keys():
                self.metric_results[metric] /= (idx + 1)
                self._update_best_metric_result(dataset_name, metric, self.metric_results[metric], current_iter)
            self._log_validation_metric_values(current_iter, dataset_name, tb_logger)

    def _initialize_best_metric_results(self, dataset_name):           # base_model.py:50-63
        if hasattr(self, 'best_metric_results') and dataset_name in self.best_metric_results:
            return
        if not hasattr(self, 'best_metric_results'):
            self.best_metric_results = dict()
        record = dict()
        for metric, content in self.opt['val']['metrics'].items():
            better = content.get('better', 'higher')
            record[metric] = dict(better=better, val=float('-inf') if better == 'higher' else float('inf'), iter=-1)
        self.best_metric_results[dataset_name] = record

    def _update_best_metric_result(self, dataset_name, metric, val, current_iter):      # base_model.py:65-73
        rec = self.best_metric_results[dataset_name][metric]
        if (rec['better'] == 'higher' and val >= rec['val']) or (rec['better'] != 'higher' and val <= rec['val']):
            rec['val'], rec['iter'] = val, current_iter

    def _log_validation_metric_values(self, current_iter, dataset_name, tb_logger):    # sr_model.py:251-264
        import logging
        log_str = f'Validation {dataset_name}\n'
        for metric, value in self.metric_results.items():
            log_str += f'\t # {metric}: {value:.4f}'
            if hasattr(self, 'best_metric_results'):
                log_str += (f'\tBest: {self.best_metric_results[dataset_name][metric]["val"]:.4f} @ '
                            f'{self.best_metric_results[dataset_name][metric]["iter"]} iter')
            log_str += '\n'
        logging.getLogger('basicsr').info(log_str)
        if tb_logger:
            for metric, value in self.metric_results.items():
                tb_logger.add_scalar(f'metrics/{dataset_name}/{metric}', value, current_iter)

    # -- convenience ---------------------------------------------------------------------------
    def test(self, lq):
        """pre_process -> (tile_)process -> post_process, as nondist_validation does per image
        (hat_model.py:136-142).  Returns the HR image (None on ranks other than dst in sharded tile mode)."""
        self.feed_data(lq)
        self.pre_process()
        if 'tile' in self.opt and self.opt['tile'] is not None:
            self.tile_process()
        else:
            self.process()
        self.post_process()
        return self.output


def _device_metric(sr_dev, gt_dev, opt_):
    """calculate_metric (basicsr/metrics/__init__.py:10-20) with the PSNR / SSIM types served on the device."""
    import copy
    o = copy.deepcopy(opt_)
    typ = o.pop('type')
    o.pop('better', None)
    if typ == 'calculate_psnr':
        return calculate_psnr(sr_dev, gt_dev, **o)
    if typ == 'calculate_ssim':
        return calculate_ssim(sr_dev, gt_dev, **o)
    try:
        from basicsr.metrics import calculate_metric
    except Exception as e:       # pragma: no cover
        raise NotImplementedError(f"metric type '{typ}' needs basicsr's METRIC_REGISTRY") from e
    data = dict(img=sr_dev.cpu().numpy())
    if gt_dev is not None:
        data['img2'] = gt_dev.cpu().numpy()
    return calculate_metric(data, opt_)


def _imwrite(img, file_path):
    """basicsr.utils.imwrite (img_util.py:149-166): makedirs + cv2.imwrite."""
    import os
    os.makedirs(os.path.abspath(os.path.dirname(file_path)), exist_ok=True)
    try:
        import cv2
    except Exception as e:       # pragma: no cover
        raise RuntimeError("saving images needs OpenCV (cv2), as in the reference") from e
    if not cv2.imwrite(file_path, img):
        raise IOError('Failed in writing images.')


try:  # the reference registers its class in basicsr's MODEL_REGISTRY (hat_model.py:13)
    from basicsr.utils.registry import MODEL_REGISTRY as _MODEL_REGISTRY
except Exception:
    _MODEL_REGISTRY = None
if _MODEL_REGISTRY is not None:
    try:
        _MODEL_REGISTRY.register(HATModel)
    except AssertionError:       # the reference's own HATModel is registered in this process: keep ours under a suffix
        _MODEL_REGISTRY.register(HATModel, suffix='b200')


def _dist_ready():
    try:
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    except Exception:
        return False
