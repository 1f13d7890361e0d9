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


def shard_tiles(rects, world_size: int):
    """Assign tiles to ranks, longest-processing-time-first on the padded input area.

    Deterministic (ties broken by the reference tile order hat_model.py:76), so every rank computes
    the same assignment without communicating.  Returns a list of `world_size` lists of tile
    indices, each in ascending (reference) order."""
    area = lambda r: (r["in_y1"] - r["in_y0"]) * (r["in_x1"] - r["in_x0"])
    order = sorted(range(len(rects)), key=lambda i: (-area(rects[i]), i))
    load = [0] * world_size
    owner = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        owner[r].append(i)
        load[r] += area(rects[i])
    return [sorted(o) for o in owner]


def group_by_shape(rects, indices):
    """Same-shape tiles -> one stacked forward.  Ordered dict {(h, w): [tile indices]}."""
    groups = OrderedDict()
    for i in indices:
        r = rects[i]
        groups.setdefault((r["in_y1"] - r["in_y0"], r["in_x1"] - r["in_x0"]), []).append(i)
    return groups


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
    peer = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)
    numel = lambda i: int(torch.Size(core_shape(rects[i], batch, channels)).numel())
    ops, bufs = [], {}
    if rank == dst:
        for r in range(world):
            if r == dst or not owners[r]:
                continue
            bufs[r] = torch.empty(sum(numel(i) for i in owners[r]), dtype=dtype, device=device)
            ops.append(dist.P2POp(dist.irecv, bufs[r], peer(r), group))
    elif owners[rank]:
        flat = torch.cat([cores[i].reshape(-1) for i in owners[rank]])
        ops.append(dist.P2POp(dist.isend, flat, peer(dst), group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    if rank != dst:
        return None
    out = dict(cores)
    for r, buf in bufs.items():
        off = 0
        for i in owners[r]:
            n = numel(i)
            out[i] = buf[off:off + n].view(core_shape(rects[i], batch, channels))
            off += n
    return out


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


class HATModel:
    """Inference half of the reference `HATModel` (hat_model.py:13-112) around a B200-native `HAT`.

    opt: the reference's option dict; the keys read are the ones the reference reads:
    `opt['network_g']['window_size']`, `opt['scale']`, and optionally `opt['tile']['tile_size']`,
    `opt['tile']['tile_pad']` (presence of `opt['tile']` selects tile mode, hat_model.py:139-141).
    group/dst: optional torch.distributed process group for tile sharding and the rank that receives the image.
    """

    def __init__(self, net, opt, group=None, dst=0):
        self.net_g = net
        self.opt = opt
        self.group = group
        self.dst = dst
        self.scale = opt.get('scale', 1)
        self.mod_pad_h = self.mod_pad_w = 0
        self.lq = self.img = self.output = None

    # -- reference API -------------------------------------------------------------------------
    def feed_data(self, data):                      # sr_model.py:87-90
        self.lq = data['lq'] if isinstance(data, dict) else data

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
        if self.group is not None or _dist_ready():
            import torch.distributed as dist
            world, rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        owners = shard_tiles(rects, world)
        self.net_g.eval()
        cores = {}
        with torch.no_grad():
            for (th, tw), idxs in group_by_shape(rects, owners[rank]).items():
                stack = torch.empty((len(idxs) * batch, channel, th, tw), dtype=img.dtype, device=img.device)
                for k, i in enumerate(idxs):        # extract (hat_model.py:77): one strided copy kernel per tile
                    r = rects[i]
                    crop_place(img, r["in_y0"], r["in_x0"], stack[k * batch:(k + 1) * batch], 0, 0, th, tw)
                out = self.net_g(stack)
                for k, i in enumerate(idxs):
                    r = rects[i]
                    core = torch.empty(core_shape(r, batch, channel), dtype=out.dtype, device=out.device)
                    crop_place(out[k * batch:(k + 1) * batch], r["crop_y0"], r["crop_x0"], core, 0, 0,
                               core.shape[2], core.shape[3])
                    cores[i] = core
        if world > 1:
            cores = gather_cores(cores, rects, owners, batch, channel, img.dtype, img.device, self.dst, self.group)
        if cores is None:                           # not the consumer rank
            self.output = None
            return
        # "start with black image" (hat_model.py:50): every output pixel is covered by exactly one core
        self.output = torch.empty((batch, channel, height * self.scale, width * self.scale), dtype=img.dtype,
                                  device=img.device)
        for i, core in cores.items():
            r = rects[i]
            crop_place(core.contiguous(), 0, 0, self.output, r["out_y0"], r["out_x0"], core.shape[2], core.shape[3])

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


def _dist_ready():
    try:
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    except Exception:
        return False
