"""TEST INFRASTRUCTURE ONLY -- generate `tests/golden/*` from the UNMODIFIED reference.

Run in the build container (where /root/reference is mounted):

    python oracle/make_golden.py

Writes, for each case in `CASES`, `tests/golden/<name>.npz` holding the reference
output for a seeded input and a seeded randomised state_dict (the generator
`oracle.hat_oracle.randomized_state_dict` is ours, so the weights can be rebuilt
on the GPU box without the reference), plus integer fixtures:
`index_tables.npz` (relative position indices straight from the reference's
buffers), `tile_rects.json` (tile geometry recovered by running the reference's
`HATModel.tile_process` with a probe network) and `state_dict_manifest.json`
(key -> shape/dtype for the five published variants).
"""
import json
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import hat_oracle as O          # noqa: E402
from oracle import load_ref                # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

BASE = dict(in_chans=3, img_size=64, window_size=16, conv_scale=0.01, overlap_ratio=0.5,
            img_range=1., mlp_ratio=2, upsampler="pixelshuffle", resi_connection="1conv")


def variant(name):
    v = dict(BASE)
    if name == "HAT-S_x4":
        v.update(embed_dim=144, depths=[6] * 6, num_heads=[6] * 6, compress_ratio=24, squeeze_factor=24, upscale=4)
    elif name in ("HAT_x2", "HAT_x3", "HAT_x4"):
        v.update(embed_dim=180, depths=[6] * 6, num_heads=[6] * 6, compress_ratio=3, squeeze_factor=30,
                 upscale=int(name[-1]))
    elif name == "HAT-L_x4":
        v.update(embed_dim=180, depths=[6] * 12, num_heads=[6] * 12, compress_ratio=3, squeeze_factor=30, upscale=4)
    else:
        raise KeyError(name)
    return v


def mini(**over):
    v = dict(BASE)
    v.update(embed_dim=48, depths=[2, 2], num_heads=[6, 6], compress_ratio=3, squeeze_factor=30, upscale=4)
    v.update(over)
    return v


# name -> (cfg, state_dict seed, input seed, input shape)
CASES = {
    "mini_x4_32x32":       (mini(), 1, 11, (1, 3, 32, 32)),
    "mini_x2_32x48":       (mini(upscale=2), 2, 12, (1, 3, 32, 48)),
    "mini_x3_48x32":       (mini(upscale=3), 3, 13, (1, 3, 48, 32)),
    "mini_x4_16x16":       (mini(), 4, 14, (1, 3, 16, 16)),
    "mini_x4_b3_32x32":    (mini(), 5, 15, (3, 3, 32, 32)),
    "mini_x2_ocabesc":     (mini(upscale=2, ocab_esc_enable=True), 6, 16, (2, 3, 32, 32)),
    "mini_x2_identity":    (mini(upscale=2, resi_connection="identity"), 7, 17, (1, 3, 32, 32)),
    "mini_x4_range255":    (mini(img_range=255.), 8, 18, (1, 3, 32, 32)),
    "hats_dims_x4_32x32":  (dict(BASE, embed_dim=144, depths=[2], num_heads=[6], compress_ratio=24,
                                 squeeze_factor=24, upscale=4), 9, 19, (1, 3, 32, 32)),
    "hat_dims_x4_32x32":   (dict(BASE, embed_dim=180, depths=[2], num_heads=[6], compress_ratio=3,
                                 squeeze_factor=30, upscale=4), 10, 20, (2, 3, 32, 32)),
    "hat_full_x4_64x64":   (variant("HAT_x4"), 21, 31, (1, 3, 64, 64)),
}


# ---- round 2: goldens at the north_star's literal bars (max-abs 1e-4 / PSNR at peak 1.0) ---------------------------
# Full-depth variants at BASELINE sizes with IMAGE-RANGE outputs: either the reference-style fresh init (seeded
# construction of OUR module, loaded into the reference; a float64 checksum of the weights is stored so a different
# torch RNG shows as a clear error) or the randomised state_dict with `conv_last.{weight,bias}` scaled by `tail_scale`
# (chosen from a 64x64 probe so that |y - mean| <= ~0.4: the dynamic-kernel / ECA / LN-affine paths stay non-zero,
# SURVEY.md 0.5, and y stays an image).  name -> (cfg, weights, sd seed, x seed, x shape, tile (size, pad) or None,
# stored subsample (row step, col step) or None).
CASES_R2 = {
    "r2_hat_x4_b16_64x64":    (variant("HAT_x4"), "random", 41, 51, (16, 3, 64, 64), None, None),     # the bench workload
    "r2_hatl_x4_64x64":       (variant("HAT-L_x4"), "random", 42, 52, (1, 3, 64, 64), None, None),
    "r2_hats_x4_64x64":       (variant("HAT-S_x4"), "random", 43, 53, (1, 3, 64, 64), None, None),
    "r2_hat_x2_64x64":        (variant("HAT_x2"), "random", 44, 54, (1, 3, 64, 64), None, None),
    "r2_hat_x3_64x64":        (variant("HAT_x3"), "random", 45, 55, (1, 3, 64, 64), None, None),
    "r2_hat_x2_256x256":      (variant("HAT_x2"), "random", 46, 56, (1, 3, 256, 256), None, None),
    "r2_hat_x3_256x256":      (variant("HAT_x3"), "random", 47, 57, (1, 3, 256, 256), None, (2, 3)),
    "r2_hat_x4_fresh_64x64":  (variant("HAT_x4"), "fresh", 0, 58, (2, 3, 64, 64), None, None),
    "r2_hatl_x4_fresh_64x64": (variant("HAT-L_x4"), "fresh", 0, 59, (1, 3, 64, 64), None, None),
    # configs[4] geometry: 4K output = 540x960 LR, reflect-padded to 544 (hat_model.py:16-26), tile 256 / pad 32
    # (12 tiles, 224- and 64-wide remainders), through the reference's own pre_process / tile_process / post_process
    "r2_hat_x4_tiled_540x960": (variant("HAT_x4"), "random", 48, 60, (1, 3, 540, 960), (256, 32), (5, 7)),
}


# ---- sibling variant HATX (reference HAT/hat/archs/hatx_arch.py) -----------------------------------
def hatx_mini(**over):
    v = dict(BASE, arch="HATX", embed_dim=48, depths=[2, 2], num_heads=[6, 6], compress_ratio=3, squeeze_factor=30,
             upscale=2, conv_scale=0.03, hab_ffn_ratio=2.0)
    v.update(over)
    return v


# name -> (cfg, state_dict seed, input seed, input shape).  The focus-bias + top-k COMBINATION is not in this list: its
# threshold falls among zero-padded keys whose scores tie at tanh(0), and torch.topk leaves the choice among equal
# scores unspecified (the oracle and the device break ties by key index; the reference differs by ~5e-4 there).
CASES_HATX = {
    "hatx_mini_x2_32x32":       (hatx_mini(), 31, 41, (2, 3, 32, 32)),
    "hatx_mini_x2_topk":        (hatx_mini(kv_topk_ratio=0.6), 32, 42, (1, 3, 32, 48)),
    "hatx_mini_x2_yaml_focus":  (hatx_mini(overlap_ratio=0.6, mlp_ratio=4.0, hab_ffn_ratio=3.0, esc_pdim=24, esc_kernel=15,
                                           ocab_esc_enable=True, ocab_esc_pdim=32, ocab_esc_kernel=17,
                                           use_focus_bias=True), 33, 43, (1, 3, 48, 32)),
    "hatx_dims_x2_32x32":       (dict(BASE, arch="HATX", embed_dim=180, depths=[2], num_heads=[6], compress_ratio=3,
                                      squeeze_factor=30, upscale=2, conv_scale=0.03, mlp_ratio=4.0, hab_ffn_ratio=3.0,
                                      use_focus_bias=True), 34, 44, (1, 3, 32, 32)),
}


def gen_hatx_cases():
    for name, (cfg, sd_seed, x_seed, shape) in CASES_HATX.items():
        torch.manual_seed(0)
        net = load_ref.build_reference_hatx(cfg)
        sd = O.randomized_state_dict(net.state_dict(), sd_seed)
        net.load_state_dict(sd, strict=True)
        x = torch.rand(shape, generator=torch.Generator().manual_seed(x_seed), dtype=torch.float32)
        y = load_ref.reference_forward(net, x)
        err = (y - O.hat_forward(x, sd, cfg)).abs().max().item()
        print(f"{name}: out {tuple(y.shape)} |y|max {y.abs().max():.3f} keys {len(sd)} oracle-vs-reference max-abs {err:.2e}")
        assert err <= 1e-4 * max(1.0, y.abs().max().item())
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), cfg=json.dumps(cfg), sd_seed=sd_seed, x_seed=x_seed,
                            x_shape=np.array(shape), y=y.numpy(), torch_version=torch.__version__)


def _load_ref_hat_model():
    load_ref.load_reference_hat()
    import importlib.util
    path = os.path.join(load_ref.REF_ROOT, "HAT", "hat", "models", "hat_model.py")
    spec = importlib.util.spec_from_file_location("hat.models.hat_model", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def reference_pipeline(net, cfg, lq, tile):
    """The reference's HATModel.pre_process -> tile_process / process -> post_process (hat_model.py:16-112) around
    `net`, driven through a stand-in `self` (the methods only touch opt / lq / net_g / scale)."""
    import contextlib
    import io
    mod = _load_ref_hat_model()

    class PerSample(torch.nn.Module):           # the eval branch of the reference is B = 1 only
        def forward(self, t):
            return load_ref.reference_forward(net, t)

    fake = types.SimpleNamespace()
    fake.opt = {"network_g": {"window_size": cfg["window_size"]}, "scale": cfg["upscale"]}
    if tile:
        fake.opt["tile"] = {"tile_size": tile[0], "tile_pad": tile[1]}
    fake.lq = lq
    fake.net_g = PerSample()
    M = mod.HATModel
    M.pre_process(fake)
    with contextlib.redirect_stdout(io.StringIO()):
        (M.tile_process if tile else M.process)(fake)
    M.post_process(fake)
    return fake.output


def gen_r2_cases(only=None):
    import time
    from super_resolution_b200.hat_arch import HAT as OurHAT
    for name, (cfg, weights, sd_seed, x_seed, shape, tile, sub) in CASES_R2.items():
        if only and name not in only:
            continue
        t0 = time.time()
        torch.manual_seed(0)
        net = load_ref.build_reference(cfg)
        extra = {}
        if weights == "fresh":
            sd = O.fresh_state_dict(OurHAT, cfg, sd_seed)
            extra["sd_checksum"] = O.state_dict_checksum(sd)
            tail_scale = 1.0
        else:
            sd = O.randomized_state_dict(net.state_dict(), sd_seed)
            probe = torch.rand((1, 3, 64, 64), generator=torch.Generator().manual_seed(x_seed), dtype=torch.float32)
            net.load_state_dict(sd, strict=True)
            y0 = load_ref.reference_forward(net, probe)
            mean = torch.tensor([0.4488, 0.4371, 0.4040]).view(1, 3, 1, 1)
            tail_scale = float(np.float32(0.4 / (y0 - mean).abs().max().item()))
            sd = O.scale_tail(sd, tail_scale)
        net.load_state_dict(sd, strict=True)
        x = torch.rand(shape, generator=torch.Generator().manual_seed(x_seed), dtype=torch.float32)
        if tile or shape[2] % cfg["window_size"] or shape[3] % cfg["window_size"]:
            y = reference_pipeline(net, cfg, x, tile)
        else:
            y = load_ref.reference_forward(net, x)
        if shape[0] * shape[2] * shape[3] <= 2 * 64 * 64:           # the restatement must agree before we commit
            err = (y - O.hat_forward(x, sd, cfg)).abs().max().item()
            assert err <= 1e-4, (name, err)
        full_shape = tuple(y.shape)
        stats = np.array([y.double().sum().item(), (y.double() ** 2).sum().item()])
        if sub:
            y = y[:, :, ::sub[0], ::sub[1]].contiguous()
        print(f"{name}: out {full_shape} stored {tuple(y.shape)} y in [{y.min():.3f}, {y.max():.3f}] "
              f"tail_scale {tail_scale:.5f}  {time.time() - t0:.0f} s", flush=True)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"),
                            cfg=json.dumps(cfg), weights=weights, sd_seed=sd_seed, x_seed=x_seed,
                            x_shape=np.array(shape), tail_scale=np.float32(tail_scale),
                            tile=np.array(tile if tile else (0, 0)), sub=np.array(sub if sub else (1, 1)),
                            y_shape=np.array(full_shape), y_stats=stats, y=y.numpy(),
                            torch_version=torch.__version__, **extra)


def gen_forward_cases():
    for name, (cfg, sd_seed, x_seed, shape) in CASES.items():
        torch.manual_seed(0)
        net = load_ref.build_reference(cfg)
        sd = O.randomized_state_dict(net.state_dict(), sd_seed)
        net.load_state_dict(sd, strict=True)
        x = torch.rand(shape, generator=torch.Generator().manual_seed(x_seed), dtype=torch.float32)
        y = load_ref.reference_forward(net, x)
        # restatement must agree before we commit the fixture
        y2 = O.hat_forward(x, sd, cfg)
        err = (y - y2).abs().max().item()
        print(f"{name}: out {tuple(y.shape)} |y|max {y.abs().max():.3f}  oracle-vs-reference max-abs {err:.2e}")
        np.savez_compressed(os.path.join(GOLD, name + ".npz"),
                            cfg=json.dumps(cfg), sd_seed=sd_seed, x_seed=x_seed,
                            x_shape=np.array(shape), y=y.numpy(), torch_version=torch.__version__)


def gen_index_tables():
    torch.manual_seed(0)
    net = load_ref.build_reference(mini())
    np.savez_compressed(os.path.join(GOLD, "index_tables.npz"),
                        rpi_sa=net.relative_position_index_SA.numpy(),
                        rpi_oca=net.relative_position_index_OCA.numpy(),
                        mask_32x48=net.calculate_mask((32, 48)).numpy())


def gen_manifest():
    man = {}
    for name in ("HAT-S_x4", "HAT_x2", "HAT_x3", "HAT_x4", "HAT-L_x4"):
        torch.manual_seed(0)
        net = load_ref.build_reference(variant(name))
        man[name] = {"cfg": variant(name),
                     "params": sum(p.numel() for p in net.parameters()),
                     "keys": {k: [list(v.shape), str(v.dtype)] for k, v in net.state_dict().items()}}
    torch.manual_seed(0)
    net = load_ref.build_reference(mini(upscale=2, ocab_esc_enable=True))
    man["mini_x2_ocabesc"] = {"cfg": mini(upscale=2, ocab_esc_enable=True),
                              "params": sum(p.numel() for p in net.parameters()),
                              "keys": {k: [list(v.shape), str(v.dtype)] for k, v in net.state_dict().items()}}
    with open(os.path.join(GOLD, "state_dict_manifest.json"), "w") as f:
        json.dump(man, f)


def gen_tile_rects():
    """Drive the reference's HATModel.tile_process with a probe network whose output
    encodes (call number, local y, local x); the input is an index image."""
    load_ref.load_reference_hat()
    import importlib.util
    path = os.path.join(load_ref.REF_ROOT, "HAT", "hat", "models", "hat_model.py")
    spec = importlib.util.spec_from_file_location("hat.models.hat_model", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = {}
    for (h, w, ts, pad, scale) in [(1024, 1024, 256, 32, 4), (544, 960, 256, 32, 4), (96, 80, 32, 16, 2),
                                   (64, 64, 64, 8, 3), (48, 112, 48, 16, 4)]:
        calls = []

        class Probe(torch.nn.Module):
            def forward(self, t):
                b, c, th, tw = t.shape
                calls.append([int(t[0, 0, 0, 0]), int(t[0, 0, -1, -1]), th, tw])
                o = torch.zeros(b, 3, th * scale, tw * scale)
                o[:, 0] = len(calls) - 1
                o[:, 1] = torch.arange(th * scale, dtype=torch.float32)[:, None]
                o[:, 2] = torch.arange(tw * scale, dtype=torch.float32)[None, :]
                return o

        fake = types.SimpleNamespace()
        fake.opt = {"tile": {"tile_size": ts, "tile_pad": pad}, "scale": scale}
        fake.scale = scale
        fake.net_g = Probe()
        idx = torch.arange(h * w, dtype=torch.float32).view(1, 1, h, w).repeat(1, 3, 1, 1)
        fake.img = idx
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            mod.HATModel.tile_process(fake)
        o = fake.output[0].to(torch.int64).numpy()
        rects = []
        for i, (first, last, th, tw) in enumerate(calls):
            ys, xs = np.nonzero(o[0] == i)
            y0, x0 = divmod(first, w)
            y1, x1 = divmod(last, w)
            rects.append(dict(idx=i, in_x0=x0, in_x1=x1 + 1, in_y0=y0, in_y1=y1 + 1,
                              out_x0=int(xs.min()), out_x1=int(xs.max()) + 1,
                              out_y0=int(ys.min()), out_y1=int(ys.max()) + 1,
                              crop_x0=int(o[2][ys.min(), xs.min()]), crop_x1=int(o[2][ys.max(), xs.max()]) + 1,
                              crop_y0=int(o[1][ys.min(), xs.min()]), crop_y1=int(o[1][ys.max(), xs.max()]) + 1))
            assert len(ys) == (rects[-1]["out_x1"] - rects[-1]["out_x0"]) * (rects[-1]["out_y1"] - rects[-1]["out_y0"])
        out[f"{h}x{w}_ts{ts}_pad{pad}_s{scale}"] = rects
        print(f"tile_rects {h}x{w} ts{ts} pad{pad}: {len(rects)} tiles")
    with open(os.path.join(GOLD, "tile_rects.json"), "w") as f:
        json.dump(out, f)


def postproc_inputs():
    """Seeded inputs of the post-processing goldens: HR-like images with out-of-range values, exact .5 ties after
    x255 (round half to even), a gray image and a pair for PSNR."""
    g = torch.Generator().manual_seed(7)
    a = torch.rand(1, 3, 40, 56, generator=g) * 1.3 - 0.15
    a[0, :, 0, :8] = torch.tensor([0.5, 1.5, 2.5, 3.5, 126.5, 127.5, 253.5, 254.5]) / 255.0     # ties
    b = (a + 0.02 * torch.randn(a.shape, generator=g)).contiguous()
    gray = torch.rand(1, 24, 32, generator=g) * 1.2 - 0.1
    wide = torch.rand(3, 16, 24, generator=g) * 300.0 - 20.0          # min_max = (0, 255)
    return a, b, gray, wide


def gen_postproc():
    """tensor2img / calculate_psnr of the unmodified reference (basicsr/utils/img_util.py, basicsr/metrics/psnr_ssim.py)."""
    load_ref.load_reference_hat()          # puts the reference's basicsr on sys.path
    from basicsr.utils.img_util import tensor2img
    from basicsr.metrics.psnr_ssim import calculate_psnr, calculate_ssim
    a, b, gray, wide = postproc_inputs()
    # the reference clamps its argument IN PLACE (`.clamp_` on a view, img_util.py:67): feed it copies
    ia, ib = tensor2img([a.clone()]), tensor2img([b.clone()])
    out = dict(a_bgr=ia, b_bgr=ib, a_rgb=tensor2img(a.clone(), rgb2bgr=False), gray=tensor2img(gray.clone()),
               wide=tensor2img(wide.clone(), min_max=(0, 255)))
    psnr = {}
    for cb in (0, 4):
        for ty in (False, True):
            psnr[f"cb{cb}_y{int(ty)}"] = float(calculate_psnr(ia, ib, crop_border=cb, test_y_channel=ty))
    psnr["gray_cb2_y1"] = float(calculate_psnr(out["gray"], tensor2img(gray * 0.97), crop_border=2, test_y_channel=True))
    ssim = {}
    for cb in (0, 4):
        for ty in (False, True):
            ssim[f"cb{cb}_y{int(ty)}"] = float(calculate_ssim(ia, ib, crop_border=cb, test_y_channel=ty))
    ssim["gray_cb2_y1"] = float(calculate_ssim(out["gray"], tensor2img(gray * 0.97), crop_border=2, test_y_channel=True))
    ssim["gray_cb0_y0"] = float(calculate_ssim(out["gray"], tensor2img(gray * 0.97), crop_border=0, test_y_channel=False))
    np.savez_compressed(os.path.join(GOLD, "postproc.npz"), psnr=json.dumps(psnr), ssim=json.dumps(ssim), **out)
    print("postproc:", {k: v.shape for k, v in out.items()}, psnr, ssim)


if __name__ == "__main__":
    if not load_ref.reference_available():
        sys.exit("reference not mounted; goldens can only be generated in the build container")
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    if len(sys.argv) > 1 and sys.argv[1] == "hatx":
        gen_hatx_cases()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "postproc":
        gen_postproc()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "r2":       # only the round-2 image-range cases (minutes of CPU)
        gen_r2_cases(set(sys.argv[2:]))
        sys.exit(0)
    gen_index_tables()
    gen_tile_rects()
    gen_postproc()
    gen_manifest()
    gen_forward_cases()
    gen_r2_cases()
    gen_hatx_cases()
