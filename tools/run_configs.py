"""Measure the BASELINE.json configurations other than the bench one (configs[2], [3], [4]) as SURVEY.md 8(d) restates
them, and print one JSON line each.  Under torchrun the tiled configurations shard their (tile, image) units over the
ranks (HATModel.tile_process) and gather the HR cores to rank 0 inside the timed region.

    python tools/run_configs.py [--cpu]              # 1 GPU (--cpu: add the CPU figure of configs[4], ~1 min)
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/run_configs.py

configs[2]  HAT SRx2 / SRx3 forward on a 256x256 LR image (bf16); parity of these is in tests/test_parity_r2.py
configs[3]  HAT-L SRx4 (12 RHAG) bf16 tiled inference of a 1024x1024 LR image (tile 256, pad 32; 16 tiles)
configs[4]  HAT SRx4 bf16, 4K OUTPUT images: 3840x2160 -> LR 540x960, reflect-padded to 544 (hat_model.py:16-26),
            tile 256 / pad 32 = 12 tiles per image with 224- and 64-wide remainders; batch sweep N = 1..32 images
Randomised weights (seed 0), uniform synthetic input; CUDA events around HATModel.test on every rank, max over ranks,
best of 3 after a warm-up (one CUDA graph per stack shape).
"""
import json
import math
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from super_resolution_b200 import HAT, HATX, HATModel  # noqa: E402

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


def make_net(cfg, cls=HAT):
    torch.manual_seed(0)
    net = cls(**cfg).eval()
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
        ms = max(e0.elapsed_time(e1), 0.0)            # device time on this rank's stream (the gather is enqueued on it)
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
    # ---- the sibling variants at benchmark scale (batch 16 of 64x64 LR, bf16) ----
    if world == 1:
        yaml_hatx = dict(conv_scale=0.03, overlap_ratio=0.6, mlp_ratio=4.0, hab_ffn_ratio=3.0, esc_pdim=24, esc_kernel=15,
                         ocab_esc_enable=True, ocab_esc_pdim=32, ocab_esc_kernel=17, kv_topk_ratio=0.6, use_focus_bias=True)
        for name, cls, cfg in (("f4 HAT SRx4 with ocab_esc_enable=True (tcgen05 attention)", HAT, base(4, ocab_esc_enable=True)),
                               ("f4 HATX SRx2, options at their defaults (spatial-gate FFN, tcgen05 attention)", HATX, base(2)),
                               ("f4 HATX SRx2, options of train_HAT_SRx2_ESC_OCAB_from_scratch.yml (ESC 24/15, OCAB ESC "
                                "32/17, overlap 0.6, focus bias, top-k 0.6: CUDA-core attention / ESC kernels)", HATX,
                                base(2, **yaml_hatx))):
            net = make_net(cfg, cls)
            x = torch.rand(16, 3, 64, 64, generator=torch.Generator().manual_seed(4)).bfloat16().to(dev)
            net(x)
            ms, y = timed(lambda: net(x))
            s_ = cfg["upscale"]
            emit(config=name, ms=round(ms, 3), out_mp_s=round(16 * 64 * 64 * s_ * s_ / 1e6 / (ms * 1e-3), 2), n_gpus=1,
                 launches=net.engine().last_launch_count, cuda_core_fallbacks=net.engine().last_fallback_count)
            del net
    # ---- configs[3]: HAT-L x4, one 1024x1024 LR image, tiled ----
    opt = {"network_g": {"window_size": 16}, "scale": 4, "tile": {"tile_size": 256, "tile_pad": 32}}
    net = make_net(base(4, 12))
    model = HATModel(net, opt, group=dist.group.WORLD if world > 1 else None, dst=0)
    x = torch.rand(1, 3, 1024, 1024, generator=torch.Generator().manual_seed(2)).bfloat16().to(dev)
    model.test(x)
    ms, y = timed(lambda: model.test(x))
    emit(config="configs[3] HAT-L SRx4 (12 RHAG) 1024x1024 LR tiled (256/32), bf16", images=1, ms=round(ms, 2),
         out_mp_s=round(4096 * 4096 / 1e6 / (ms * 1e-3), 2), n_gpus=world, tiles_per_image=16,
         out_shape=list(y.shape) if y is not None else None)
    del net, model
    # ---- configs[4]: HAT x4, 4K-output images (540x960 LR), batch sweep ----
    net = make_net(base(4))
    model = HATModel(net, opt, group=dist.group.WORLD if world > 1 else None, dst=0)
    for nimg in (1, 2, 4, 8, 16, 32):
        x = torch.rand(nimg, 3, 540, 960, generator=torch.Generator().manual_seed(3)).bfloat16().to(dev)
        model.test(x)
        ms, y = timed(lambda: model.test(x))
        emit(config="configs[4] HAT SRx4 4K output (3840x2160; LR 540x960 reflect-padded to 544) tiled (256/32), bf16",
             images=nimg, ms=round(ms, 2), ms_per_image=round(ms / nimg, 2),
             out_mp_s=round(nimg * 3840 * 2160 / 1e6 / (ms * 1e-3), 2), n_gpus=world, tiles_per_image=12,
             out_shape=list(y.shape) if y is not None else None)
        del x, y
    del net, model
    # ---- CPU figure for configs[4]: the oracle port of the reference forward on the host cores ----
    if rank == 0 and "--cpu" in sys.argv:
        from oracle import hat_oracle as O
        cfg = base(4)
        torch.manual_seed(0)
        sd = {k: v.clone() for k, v in HAT(**cfg).state_dict().items()}
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        # one small tile class timed, the image scaled by processed LR pixels (the forward is local: cost ~ pixels)
        th, tw = 64, 224
        t = torch.rand(1, 3, th, tw)
        O.hat_forward(t[:, :, :16, :16], sd, cfg)
        t0 = time.perf_counter()
        O.hat_forward(t, sd, cfg)
        sec = time.perf_counter() - t0
        px_image = 774144                       # processed LR pixels per image incl. tile padding
        sec_image = sec * px_image / (th * tw)
        emit(config="configs[4] CPU: oracle port of the reference forward, fp32", cores=threads,
             sample=f"one {th}x{tw} tile timed ({sec:.1f} s), scaled by processed LR pixels to the 12 tiles of an image",
             s_per_image_scaled=round(sec_image, 1), out_mp_s=round(3840 * 2160 / 1e6 / sec_image, 4))
if world > 1:
    dist.destroy_process_group()
