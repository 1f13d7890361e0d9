"""Tile-mode caller (super_resolution_b200.hat_model): planning + the one collective on CPU (gloo, world 2),
and the device pipeline against the oracle's restatement of hat_model.py:16-112 on the GPU."""
import json
import os
import socket

import pytest
import torch

from oracle import hat_oracle as O
from super_resolution_b200 import hat_model as HM
from tests.helpers import GOLDEN


def _rects(h, w, ts, tp, s):
    return O.tile_rects(h, w, ts, tp, s)


@pytest.mark.parametrize("geom", [(1024, 1024, 256, 32, 4), (544, 960, 256, 32, 4), (64, 64, 32, 16, 2),
                                  (48, 80, 32, 16, 3)])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_shard_tiles_partitions_every_tile_once_and_balances(geom, world):
    h, w, ts, tp, s = geom
    rects = _rects(h, w, ts, tp, s)
    owners = HM.shard_tiles(rects, world)
    flat = sorted(i for o in owners for i in o)
    assert flat == list(range(len(rects)))                       # a partition
    assert owners == HM.shard_tiles(rects, world)                # deterministic
    area = lambda i: (rects[i]["in_y1"] - rects[i]["in_y0"]) * (rects[i]["in_x1"] - rects[i]["in_x0"])
    loads = [sum(area(i) for i in o) for o in owners]
    if len(rects) >= world:
        assert max(loads) - min(loads) <= max(area(i) for i in range(len(rects)))   # LPT bound
    # every HR pixel is covered by exactly one core
    cover = torch.zeros(h * s, w * s, dtype=torch.int32)
    for r in rects:
        cover[r["out_y0"]:r["out_y1"], r["out_x0"]:r["out_x1"]] += 1
    assert int(cover.min()) == 1 and int(cover.max()) == 1


def test_config4_assignment_is_the_expected_mix():
    """BASELINE configs[3]: 1024x1024 LR, tile 256 / pad 32 -> 4 corner 288x288, 8 edge, 4 interior 320x320."""
    rects = _rects(1024, 1024, 256, 32, 4)
    shapes = sorted((r["in_y1"] - r["in_y0"], r["in_x1"] - r["in_x0"]) for r in rects)
    assert shapes.count((288, 288)) == 4 and shapes.count((320, 320)) == 4 and len(shapes) == 16
    for world in (2, 4, 8):
        owners = HM.shard_tiles(rects, world)
        assert all(len(o) == 16 // world for o in owners)
    groups = HM.group_by_shape(rects, list(range(16)))
    assert sum(len(v) for v in groups.values()) == 16 and len(groups) == 4


def test_shards_use_the_committed_reference_geometry():
    """The plan the ranks share is the golden geometry generated from the unmodified reference."""
    from super_resolution_b200 import _lib
    with open(os.path.join(GOLDEN, "tile_rects.json")) as f:
        gold = json.load(f)
    for key, rects in gold.items():
        hw, ts, pad, s = key.split("_")
        h, w = (int(v) for v in hw.split("x"))
        got = _lib.tile_plan(h, w, int(ts[2:]), int(pad[3:]), int(s[1:]))
        assert got == rects, key


def test_mod_pad_matches_oracle():
    for h, w in [(540, 960), (64, 64), (37, 45), (16, 17)]:
        assert HM.mod_pad(h, w, 16) == O.mod_pad(h, w, 16)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_split_stacks_caps_the_stacked_forward():
    """2048x2048 LR at tile 256 / pad 32 has 36 interior 320x320 tiles: one stack of 36 would exceed the 2^21-token
    limit of hatb200_forward; the plan cuts it into forwards of at most max_tokens LR pixels."""
    rects = O.tile_rects(2048, 2048, 256, 32, 4)
    groups = HM.group_by_shape(rects, list(range(len(rects))))
    assert len(groups[(320, 320)]) == 36
    seen = []
    for (th, tw), idxs in groups.items():
        parts = HM.split_stacks(idxs, th, tw, 1)
        assert [i for p in parts for i in p] == idxs                       # order kept, nothing lost
        for p in parts:
            assert len(p) * th * tw <= HM.DEFAULT_STACK_TOKENS < HM.MAX_FORWARD_TOKENS
        seen += idxs
    assert sorted(seen) == list(range(64))
    assert [len(p) for p in HM.split_stacks(groups[(320, 320)], 320, 320, 1, 1 << 20)] == [10, 10, 10, 6]
    assert [len(p) for p in HM.split_stacks(groups[(320, 320)], 320, 320, 1, 1 << 18)] == [2] * 18
    # a single tile that is larger than the cap still gets its own forward (the C library then checks the hard limit)
    assert HM.split_stacks([0, 1], 1024, 1024, 2, max_tokens=1 << 20) == [[0], [1]]
    # 2560x2560 at 512/32 (576x576 interior tiles, the shipped HAT_tile_example geometry): 3 per stack
    assert [len(p) for p in HM.split_stacks(list(range(9)), 576, 576, 1, 1 << 20)] == [3, 3, 3]


def _rank0_only_worker(rank, world, port, q):
    """The reference validates on rank 0 only (sr_model.py:180-182): with an initialised default group and no
    group / shard option, tile_process must not wait for the other ranks."""
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        if rank == 0:
            calls = []

            class Net:
                def eval(self):
                    return self

                def __call__(self, t):
                    calls.append(tuple(t.shape))
                    return t.repeat_interleave(2, 2).repeat_interleave(2, 3)

            m = HM.HATModel(Net(), {"network_g": {"window_size": 16}, "scale": 2, "tile": {"tile_size": 32, "tile_pad": 16}})
            m.img = torch.zeros(1, 3, 64, 64)
            # stop at the device boundary: the planning above it is what is under test
            try:
                m.tile_process()
                q.put("ran without a CUDA image")
            except RuntimeError as e:
                q.put("no CPU path" in str(e))
        else:
            q.put(True)                         # rank 1 never calls tile_process
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_tile_process_does_not_shard_without_group_or_flag():
    import inspect
    src = inspect.getsource(HM.HATModel.tile_process)
    assert "self.group is not None or (self.opt['tile'].get('shard')" in src
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank0_only_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r is True for r in res), res


# ---- world_size-2 gather over gloo (host logic of the N>1 path) ---------------------------------------
def _worker(rank, world, port, geom, q):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        h, w, ts, tp, s = geom
        batch, ch = 2, 3
        rects = O.tile_rects(h, w, ts, tp, s)
        owners = HM.shard_tiles(rects, world)
        cores = {}
        for i in owners[rank]:                       # a fake "forward": the value encodes tile, sample and position
            shp = HM.core_shape(rects[i], batch, ch)
            cores[i] = (torch.arange(torch.Size(shp).numel(), dtype=torch.float32).view(shp) * 1e-3 + i)
        got = HM.gather_cores(cores, rects, owners, batch, ch, torch.float32, torch.device("cpu"), dst=0)
        if rank == 0:
            ok = sorted(got) == list(range(len(rects)))
            for i, t in got.items():
                shp = HM.core_shape(rects[i], batch, ch)
                want = torch.arange(torch.Size(shp).numel(), dtype=torch.float32).view(shp) * 1e-3 + i
                ok = ok and torch.equal(t, want)
            q.put(bool(ok))
        else:
            q.put(got is None)
    finally:
        dist.destroy_process_group()


def _unit_worker(rank, world, port, geom, batch, q):
    """tile_process's own exchange (units = one tile of one sample, flat send buffer, exchange_flat, flat_views)."""
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        h, w, ts, tp, s = geom
        rects = O.tile_rects(h, w, ts, tp, s)
        units = [dict(rects[i], tile=i, sample=b) for i in range(len(rects)) for b in range(batch)]
        owners = HM.shard_tiles(units, world)
        assert sorted(sum(owners, [])) == list(range(len(units)))
        val = lambda u: torch.arange(torch.Size(HM.core_shape(units[u], 1, 3)).numel(), dtype=torch.float32).view(
            HM.core_shape(units[u], 1, 3)) * 1e-3 + u
        send = None
        if rank != 0:
            send = torch.empty(HM.flat_numel(units, owners[rank], 1, 3))
            for u, view in HM.flat_views(send, units, owners[rank], 1, 3).items():
                view.copy_(val(u))
        bufs = HM.exchange_flat(send, units, owners, 1, 3, torch.float32, torch.device("cpu"), dst=0)
        if rank == 0:
            ok = sorted(bufs) == [r for r in range(1, world) if owners[r]]
            for r, buf in bufs.items():
                for u, core in HM.flat_views(buf, units, owners[r], 1, 3).items():
                    ok = ok and torch.equal(core, val(u))
            q.put(bool(ok))
        else:
            q.put(bufs == {})
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("geom,batch", [((96, 128, 32, 16, 2), 1), ((64, 80, 32, 16, 4), 3), ((544, 960, 256, 32, 4), 4)])
def test_unit_exchange_world2_gloo(geom, batch):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_unit_worker, args=(r, 2, port, geom, batch, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(res)


def test_shape_aware_plans():
    """configs[3] at 8 ranks: every rank gets a same-shape pair (one stacked forward); a batch of 4K-output images
    (more than 32 units) is partitioned shape-major so that a rank holds few shapes."""
    rects = O.tile_rects(1024, 1024, 256, 32, 4)
    shape = lambda r: (r["in_y1"] - r["in_y0"], r["in_x1"] - r["in_x0"])
    owners = HM.shard_tiles(rects, 8)
    assert all(len(o) == 2 and shape(rects[o[0]]) == shape(rects[o[1]]) for o in owners)
    rects = O.tile_rects(544, 960, 256, 32, 4)
    units = [dict(rects[i], tile=i, sample=b) for i in range(len(rects)) for b in range(8)]
    owners = HM.shard_tiles(units, 8)
    assert sorted(sum(owners, [])) == list(range(96))
    area = lambda o: sum(shape(units[i])[0] * shape(units[i])[1] for i in o)
    assert max(area(o) for o in owners) <= 1.1 * sum(area(o) for o in owners) / 8
    assert max(len({shape(units[i]) for i in o}) for o in owners) <= 4
    assert owners == HM.shard_tiles(units, 8)


@pytest.mark.parametrize("geom", [(96, 128, 32, 16, 2), (64, 64, 64, 16, 4)])
def test_gather_cores_world2_gloo(geom):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, geom, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(res)


# ---- device pipeline vs the oracle ---------------------------------------------------------------
MINI = dict(upscale=2, in_chans=3, img_size=64, window_size=16, compress_ratio=3, squeeze_factor=30, conv_scale=0.01,
            overlap_ratio=0.5, img_range=1., depths=[2, 2], embed_dim=48, num_heads=[6, 6], mlp_ratio=2,
            upsampler='pixelshuffle', resi_connection='1conv')


def _mini_net(scale):
    from super_resolution_b200 import HAT
    cfg = dict(MINI, upscale=scale)
    net = HAT(**cfg).eval()
    sd = O.randomized_state_dict(net.state_dict(), 3)
    net.load_state_dict(sd, strict=True)
    return cfg, sd, net.to("cuda:0")


@pytest.mark.gpu
@pytest.mark.parametrize("scale,hw,tile", [(2, (40, 70), (32, 16)), (4, (48, 48), (32, 16)), (3, (33, 64), (32, 16))])
def test_tile_mode_matches_reference_pipeline_fp32(scale, hw, tile):
    cfg, sd, net = _mini_net(scale)
    x = torch.rand(2, 3, *hw, generator=torch.Generator().manual_seed(5))
    want = O.tiled_forward(x, sd, cfg, tile[0], tile[1])
    opt = {"network_g": {"window_size": 16}, "scale": scale, "tile": {"tile_size": tile[0], "tile_pad": tile[1]}}
    got = HM.HATModel(net, opt).test(x.to("cuda:0")).float().cpu()
    assert got.shape == want.shape
    assert (got - want).abs().max().item() <= 1e-4 * max(1.0, want.abs().max().item())


@pytest.mark.gpu
def test_full_image_mode_pads_and_crops_like_the_reference():
    cfg, sd, net = _mini_net(2)
    x = torch.rand(1, 3, 37, 45, generator=torch.Generator().manual_seed(6))
    ph, pw = O.mod_pad(37, 45, 16)
    xp = torch.nn.functional.pad(x, (0, pw, 0, ph), "reflect")
    want = O.hat_forward(xp, sd, cfg)[:, :, :37 * 2, :45 * 2]
    opt = {"network_g": {"window_size": 16}, "scale": 2}
    got = HM.HATModel(net, opt).test(x.to("cuda:0")).float().cpu()
    assert got.shape == want.shape
    assert (got - want).abs().max().item() <= 1e-4 * max(1.0, want.abs().max().item())


@pytest.mark.gpu
def test_tile_mode_bf16_psnr():
    from tests.helpers import psnr
    cfg, sd, net = _mini_net(4)
    x = torch.rand(1, 3, 64, 80, generator=torch.Generator().manual_seed(7))
    want = O.tiled_forward(x, sd, cfg, 32, 16)
    opt = {"network_g": {"window_size": 16}, "scale": 4, "tile": {"tile_size": 32, "tile_pad": 16}}
    got = HM.HATModel(net, opt).test(x.to("cuda:0").bfloat16()).float().cpu()
    assert psnr(got, want, peak=max(1.0, want.abs().max().item())) >= 50.0


def test_no_cpu_path():
    from super_resolution_b200 import HAT
    net = HAT(**MINI).eval()
    opt = {"network_g": {"window_size": 16}, "scale": 2, "tile": {"tile_size": 32, "tile_pad": 16}}
    with pytest.raises(RuntimeError):
        HM.HATModel(net, opt).test(torch.rand(1, 3, 40, 40))
