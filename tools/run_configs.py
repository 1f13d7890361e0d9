"""Measure the BASELINE.json configurations other than the bench one (configs[2], [3], [4]) and print one JSON line
each.  Under torchrun the tiled configurations shard their tiles over the ranks (HATModel.tile_process).

    python tools/run_configs.py                       # 1 GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/run_configs.py

configs[2]  HAT SRx2 / SRx3 forward on a 256x256 LR image (bf16), PSNR of the bf16 path against our fp32 path
configs[3]  HAT-L SRx4 (12 RHAG) bf16 tiled inference of a 1024x1024 LR image (tile 256, pad 32)
configs[4]  HAT SRx4 tiled inference of 1024x1024 LR images (4096x4096 = "4K" output), batch sweep
Random-init weights (randomised state_dict, seed 0), uniform synthetic input; timing with CUDA events, best of 3.
"""
import json
import math
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from super_resolution_b200 import HAT, HATModel  # noqa: E402

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def base(scale, depths=6, **over):
    cfg = dict(upscale=scale, in_chans=3, img_size=64, window_size=16, compress_ratio=3, squeeze_factor=30,
               conv_scale=0.01, overlap_ratio=0.5, img_range=1., depths=[6] * depths, embed_dim=180,
               num_heads=[6] * depths, mlp_ratio=2, upsampler='pixelshuffle', resi_connection='1conv')
    cfg.update(over)
    return cfg


def make_net(cfg):
    torch.manual_seed(0)
    net = HAT(**cfg).eval()
    sd = net.state_dict()
    g = torch.Generator().manual_seed(0)
    for k, v in sd.items():      # a fresh init leaves the dynamic-kernel branch at zero and every LayerNorm at identity:
        if not v.is_floating_point():      # redraw like the parity tests do (weights N(0,.05), biases N(0,.02), LN 1+N(0,.1))
            continue
        r = torch.randn(v.shape, generator=g)
        if "norm" in k.split(".")[-2] and k.endswith(".weight"):
            sd[k] = 1.0 + 0.1 * r
        elif k.endswith("relative_position_bias_table") or "dwc_proj" in k:
            sd[k] = 0.2 * r
        elif k.endswith(".bias") or k.endswith("plk_filter") or k.endswith("esc_plk"):
            sd[k] = 0.02 * r
        else:
            sd[k] = 0.05 * r
    net.load_state_dict(sd)
    return net.to(dev)


def timed(fn, reps=3):
    best = float("inf")
    out = None
    for _ in range(reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = max(e0.elapsed_time(e1), 0.0)
        wall = (time.perf_counter() - t0) * 1e3
        ms = max(ms, wall) if world > 1 else ms        # sharded: host-side gather included
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        best = min(best, ms)
    return best, out


def emit(**kw):
    if rank == 0:
        print(json.dumps(kw), flush=True)


def psnr(a, b, peak=1.0):
    mse = torch.mean((a.float() - b.float()) ** 2).item()
    return 10 * math.log10(peak * peak / max(mse, 1e-30))


with torch.no_grad():
    # ---- configs[2]: x2 and x3 on 256x256 ----
    if world == 1:
        for s in (2, 3):
            net = make_net(base(s))
            x = torch.rand(1, 3, 256, 256, generator=torch.Generator().manual_seed(1)).to(dev)
            y32 = net(x)
            net(x.bfloat16())
            ms, y16 = timed(lambda: net(x.bfloat16()))
            emit(config=f"configs[2] HAT SRx{s} 256x256 LR, batch 1, bf16", ms=round(ms, 3),
                 out_mp_s=round(256 * 256 * s * s / 1e6 / (ms * 1e-3), 2), psnr_bf16_vs_fp32_db=round(psnr(y16, y32, max(1.0, y32.abs().max().item())), 2),
                 n_gpus=1)
            del net
    # ---- configs[3]: HAT-L x4, 1024x1024 tiled; configs[4]: HAT x4, 4K output tiled ----
    for name, cfg, nimg in (("configs[3] HAT-L SRx4 (12 RHAG) 1024x1024 LR tiled (256/32), bf16", base(4, 12), 1),
                            ("configs[4] HAT SRx4 1024x1024 LR -> 4096x4096 tiled (256/32), bf16", base(4), 1),
                            ("configs[4] HAT SRx4 1024x1024 LR -> 4096x4096 tiled (256/32), bf16", base(4), 2)):
        net = make_net(cfg)
        opt = {"network_g": {"window_size": 16}, "scale": 4, "tile": {"tile_size": 256, "tile_pad": 32}}
        model = HATModel(net, opt, group=dist.group.WORLD if world > 1 else None, dst=0)
        x = torch.rand(nimg, 3, 1024, 1024, generator=torch.Generator().manual_seed(2)).bfloat16().to(dev)
        model.test(x)            # warm-up: graphs per tile shape
        ms, y = timed(lambda: model.test(x), reps=2)
        emit(config=name, images=nimg, ms=round(ms, 2), out_mp_s=round(nimg * 4096 * 4096 / 1e6 / (ms * 1e-3), 2),
             n_gpus=world, tiles_per_image=16, out_shape=list(y.shape) if y is not None else None)
        del net, model
if world > 1:
    dist.destroy_process_group()
