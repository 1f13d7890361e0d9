#!/usr/bin/env python
"""bench.py -- HAT-SRx4 output megapixels/s on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): HAT SRx4 (embed_dim 180, 6x6 RHAG) bf16 forward, batch 16 of
64x64 synthetic LR patches per GPU, random-init weights (seed 0), uniform[0,1) input (seed 0).
A step = one forward over one batch.  Weak scaling: every rank runs its own batch (image-batch
sharding), no collective on the data path; value = all ranks' output pixels / max-over-
ranks device time.  One JSON line on rank 0.

--impl reference: the reference's CPU forward (the oracle port -- the Python reference itself cannot
travel to the GPU box) on the host cores, each step a bounded sample of the same workload.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

HAT_X4 = dict(upscale=4, in_chans=3, img_size=64, window_size=16, compress_ratio=3, squeeze_factor=30,
              conv_scale=0.01, overlap_ratio=0.5, img_range=1., depths=[6] * 6, embed_dim=180,
              num_heads=[6] * 6, mlp_ratio=2, upsampler='pixelshuffle', resi_connection='1conv')
BATCH, LR = 16, 64
MFLOP_PER_LR_PX = 45.3464          # SURVEY.md 8d / BASELINE.md 2 (algorithmic, no padding)
METRIC = "HAT-SRx4 output megapixels/s (bf16 forward, batch 16 of 64x64 LR per GPU)"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tf=d["bf16_tflops_sustained"], tf_burst=d["bf16_tflops"], src="measured")
    return dict(hbm=6650.0, tf=1400.0, tf_burst=1590.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark(self):
        """Start of the timed region: only samples taken from here on count (the sampler itself is started before the
        warm-up, nvidia-smi needs 0.1-1 s to come up -- longer with eight ranks starting one each)."""
        self.first = len(self.lines)

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.03)
        self.proc.terminate()
        window = "timed region"
        if len(self.lines) > getattr(self, "first", 0):
            self.lines = self.lines[getattr(self, "first", 0):]
        else:                                   # the region was shorter than the sampler's first answer: warm-up samples
            window = "warm-up (same workload) + timed region"
        self.window = window
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [v.strip() for v in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": self.window}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def cpu_forward_sample(n_samples, threads):
    """Oracle port of the reference forward on the host cores: seconds for n_samples 64x64 patches."""
    from oracle import hat_oracle as O
    from super_resolution_b200 import HAT
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    net = HAT(**HAT_X4)
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    x = torch.rand(n_samples, 3, LR, LR, generator=torch.Generator().manual_seed(0))
    O.hat_forward(x[:1], sd, HAT_X4)            # warm-up
    t0 = time.perf_counter()
    O.hat_forward(x, sd, HAT_X4)
    return time.perf_counter() - t0


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n = 2                                     # LR patches per step (bounded sample of the batch of 16)
    from oracle import hat_oracle as O
    from super_resolution_b200 import HAT
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    net = HAT(**HAT_X4)
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    x = torch.rand(n, 3, LR, LR, generator=torch.Generator().manual_seed(0))
    for _ in range(max(1, min(args.warmup, 2))):
        O.hat_forward(x[:1], sd, HAT_X4)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        O.hat_forward(x, sd, HAT_X4)
        times.append(time.perf_counter() - t0)
    t = sum(times) / len(times)
    out_mp = n * (LR * 4) ** 2 / 1e6
    val = out_mp / t
    sample = (f"{n} of the {BATCH} 64x64 LR patches per step, fp32, oracle port of the reference forward "
              f"(torch-CPU ops, {threads} threads); scaled figures assume linear cost in patches")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "MP/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "HAT SRx4 (embed_dim=180, 6x6 RHAG) forward, 64x64 LR patches, CPU sample",
                   "patches_per_step": n, "full_batch": BATCH, "ms_per_full_batch_scaled": t * 1e3 * BATCH / n},
        "cpu_baseline": {"value": val, "unit": "MP/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "MP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# name fragments of each kernel family in an ncu launch list (profiles/r*_ncu_full_*_summary.csv)
NCU_NAMES = {"gemm_tc": ("gemm_tc_kernel",), "conv3_tc": ("conv3_tc_kernel",), "tapconv_tc": ("tapconv_tc_kernel",),
             "plk_tc": ("plk_tc_kernel",), "attn_tc": ("attn_tc_kernel",), "dwgate": ("dwgate",),
             "ffn_tc": ("ffn_tc_kernel",), "layernorm": ("layernorm_kernel",)}


def ncu_traffic(name_parts):
    """Mean dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes) of the kernels whose name contains one of
    `name_parts`, from the newest profiles/r*_ncu_full_*_summary.csv; (None, None) when there is no capture."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_*_summary.csv")))
    if not files:
        return None, None
    try:
        rows = list(csv.reader(open(files[-1])))
        hdr, units = rows[0], rows[1]
        col = lambda suffix: next(i for i, h in enumerate(hdr) if h.endswith(suffix))
        ir, iw = col("dram__bytes_read.sum"), col("dram__bytes_write.sum")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        vals = [float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
                for r in rows[2:] if any(n in r[0] for n in name_parts)]
        if not vals:
            return None, None
        return sum(vals) / len(vals), f"profiles/{os.path.basename(files[-1])} ({len(vals)} launches, ncu --set full)"
    except Exception:      # a malformed capture must not take the bench down
        return None, None


def summarize_profile(prof, peaks, tc_on):
    """Group the per-op profile by the KERNEL that served each op (hatb200_op_profile names it), find the dominant
    kernel by measured time inside the step, and state ITS roofline: the bound is whichever of (algorithmic flops /
    bf16 peak) and (algorithmic bytes / copy bandwidth) is the larger time."""
    groups = {}
    for tag, r in prof.items():
        g = groups.setdefault(r.get("kernel", "?"), dict(launches=0, ms=0.0, flops=0.0, bytes=0.0, tags=[]))
        for f in ("launches", "ms", "flops", "bytes"):
            g[f] += r[f]
        g["tags"].append(tag)
    total_ms = sum(g["ms"] for g in groups.values())

    def roof_of(g):
        t_tensor = g["flops"] / (peaks["tf"] * 1e12)
        t_hbm = g["bytes"] / (peaks["hbm"] * 1e9)
        sec = g["ms"] * 1e-3
        if t_tensor >= t_hbm:
            ach = g["flops"] / sec / 1e12
            return {"bound": "tensor", "achieved": ach, "peak": peaks["tf"], "unit": "TFLOP/s", "frac": ach / peaks["tf"]}
        ach = g["bytes"] / sec / 1e9
        return {"bound": "hbm", "achieved": ach, "peak": peaks["hbm"], "unit": "GB/s", "frac": ach / peaks["hbm"]}

    dom = max(groups, key=lambda k: groups[k]["ms"])
    g = groups[dom]
    roof = roof_of(g)
    # DRAM traffic per launch of the dominant kernel from the committed `ncu --set full` capture (cold cache: ncu
    # flushes L2 between replays), next to the algorithmic bytes per launch of the same kernel
    traffic, traffic_src = ncu_traffic(NCU_NAMES.get(dom, (dom,)))
    roof["traffic"] = traffic
    roof["traffic_source"] = traffic_src
    roof["algorithmic_bytes_per_launch"] = g["bytes"] / max(1, g["launches"])
    roof["algorithmic_flops_per_launch"] = g["flops"] / max(1, g["launches"])
    roof.update({"kernel": dom, "ops": g["tags"], "launches": g["launches"],
                 "avg_launch_us": g["ms"] * 1e3 / max(1, g["launches"]),
                 "share_of_step": g["ms"] / total_ms if total_ms else None,
                 "peak_source": f"{peaks['src']} (sustained bf16 / copy bandwidth, MEASURED_PEAKS.json)"})
    per_kernel = {}
    for k, gk in sorted(groups.items(), key=lambda kv: -kv[1]["ms"]):
        rk = roof_of(gk)
        per_kernel[k] = {"launches": gk["launches"], "ms": round(gk["ms"], 4), "bound": rk["bound"],
                         "achieved": round(rk["achieved"], 1), "unit": rk["unit"], "frac": round(rk["frac"], 4),
                         "share_of_step": round(gk["ms"] / total_ms, 4) if total_ms else None}
    per_op = {t: {"kernel": r.get("kernel"), "launches": r["launches"], "ms": round(r["ms"], 4),
                  "tflops": round(r["flops"] / (r["ms"] * 1e-3) / 1e12, 2) if r["ms"] > 0 else None,
                  "gbs": round(r["bytes"] / (r["ms"] * 1e-3) / 1e9, 1) if r["ms"] > 0 else None}
              for t, r in prof.items()}
    return roof, per_kernel, per_op, total_ms


HATL_X4 = dict(HAT_X4, depths=[6] * 12, num_heads=[6] * 12)


def run_tiled(world, rank, dev, runs=3):
    """BASELINE configs[3] inside a timed region: HAT-L SRx4 (12 RHAG) bf16, ONE 1024x1024 LR image, tile 256 / pad 32
    (16 tiles: 4 of 288x288, 8 of 288x320, 4 of 320x320), tiles sharded over the ranks longest-processing-time-first,
    cropped HR cores gathered to rank 0 (NCCL send / recv over NVLink) INSIDE the timed region.  Strong scaling of one
    image: CUDA events around HATModel.test on every rank, max over ranks; rank 0 also runs the same job alone so that
    the efficiency against one GPU comes from the same process and box."""
    import torch.distributed as dist
    from super_resolution_b200 import HAT, HATModel
    torch.manual_seed(0)
    net = HAT(**HATL_X4).eval().to(dev)
    x = torch.rand(1, 3, 1024, 1024, generator=torch.Generator().manual_seed(0)).bfloat16().to(dev)
    opt = {"network_g": {"window_size": 16}, "scale": 4, "tile": {"tile_size": 256, "tile_pad": 32}}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(model):
        model.record_events = True
        for _ in range(2):                    # warm-up: one CUDA graph per stack shape
            model.test(x)
        tot, comp = [], []
        for _ in range(runs):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            model.test(x)
            e1.record()
            torch.cuda.synchronize()
            tot.append(e0.elapsed_time(e1))
            comp.append(e0.elapsed_time(model.events['compute_end']))
        return sum(tot) / runs, sum(comp) / runs

    out_mp = 4096 * 4096 / 1e6
    res = {"config": "BASELINE configs[3]: HAT-L SRx4 (12 RHAG) bf16, one 1024x1024 LR image, tile 256 / pad 32, "
                     "16 tiles sharded over the ranks (LPT), HR cores gathered to rank 0 inside the timed region",
           "n_gpus": world, "tiles": 16, "runs": runs, "scaling": "strong"}
    t1 = None
    if rank == 0:                             # the same job on one GPU (the other ranks wait at the barrier below)
        world_, world = world, 1
        t1, _ = timed(HATModel(net, opt))
        world = world_
    if world == 1:
        res.update({"ms": t1, "out_mp_s": out_mp / (t1 * 1e-3)})
        return res
    dist.barrier()
    model = HATModel(net, opt, group=dist.group.WORLD, dst=0)
    t, c = timed(model)
    v = torch.tensor([t, c, c], device=dev, dtype=torch.float64)
    mx = v.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = v.clone()
    dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    t_max, c_max, c_mean = mx[0].item(), mx[1].item(), sm[2].item() / world
    if rank == 0:
        res.update({"ms": t_max, "out_mp_s": out_mp / (t_max * 1e-3), "ms_1gpu": t1,
                    "efficiency_vs_1gpu": t1 / (world * t_max), "compute_ms_max": c_max, "compute_ms_mean": c_mean,
                    "imbalance": c_max / c_mean, "gather_ms": t_max - c_max,
                    "gather_bytes": int(4096 * 4096 * 3 * 2 * (world - 1) / world)})
    return res


def run_ours(args):
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        sys.exit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU arm)")
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from super_resolution_b200 import HAT
    peaks = load_peaks()
    torch.manual_seed(0)
    net = HAT(**HAT_X4).eval().to(dev)
    eng = net.engine()
    tc_on = os.environ.get("HATB200_TENSOR_CORES", "1") != "0"
    eng.set_option("tensor_cores", int(tc_on))
    g = torch.Generator().manual_seed(rank)
    x_host = torch.rand(BATCH, 3, LR, LR, generator=g).bfloat16().pin_memory()
    x = x_host.to(dev)
    out_px = BATCH * (LR * 4) ** 2                           # delivered HR pixels per rank per step
    W, K = max(3, args.warmup), args.steps
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    clocks = ClockSampler(local)
    clocks.start()                                           # comes up during the warm-up; windowed by mark() below
    for _ in range(W):
        y = net(x)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput: K steps, CUDA events per step, L2 flushed between steps ----
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    clocks.mark()
    for e0, e1 in ev:
        flush.zero_()
        e0.record()
        y = net(x)
        e1.record()
    barrier()
    clk = clocks.stop()
    total_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev)
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = t.item() / K
    value = world * out_px / 1e6 / (ms_step * 1e-3)

    # ---- end to end through the public API: pinned host input -> H2D -> net(x) -> D2H of the HR batch ----
    y_host = torch.empty((BATCH, 3, LR * 4, LR * 4), dtype=torch.bfloat16).pin_memory()
    xd = torch.empty_like(x)
    for _ in range(2):
        xd.copy_(x_host, non_blocking=True)
        y_host.copy_(net(xd), non_blocking=True)
    barrier()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(K):
        xd.copy_(x_host, non_blocking=True)
        y_host.copy_(net(xd), non_blocking=True)
    s1.record()
    barrier()
    t2 = torch.tensor([s0.elapsed_time(s1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    e2e_val = world * out_px / 1e6 / (t2.item() / K * 1e-3)

    launches = eng.last_launch_count
    tiled = None
    if not args.no_tiled:
        del flush
        torch.cuda.empty_cache()
        tiled = run_tiled(world, rank, dev)
    if rank == 0:
        prof = eng.op_profile(x)
        torch.cuda.synchronize()
        roof, per_kernel, per_op, prof_ms = summarize_profile(prof, peaks, tc_on)
        lr_px = BATCH * LR * LR
        whole = MFLOP_PER_LR_PX * 1e6 * lr_px / (ms_step * 1e-3) / 1e12
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            t1 = cpu_forward_sample(1, threads)
            n = int(min(BATCH, max(2, 12.0 / max(t1, 1e-3))))
            tn = cpu_forward_sample(n, threads)
            cpu = {"value": n * (LR * 4) ** 2 / 1e6 / tn, "unit": "MP/s", "cores": threads, "kind": "port",
                   "sample": f"{n} of the {BATCH} 64x64 LR patches, fp32, oracle port of the reference forward on "
                             f"{threads} host threads ({tn:.2f} s)"}
        line = {
            "metric": METRIC, "value": value, "unit": "MP/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "HAT SRx4 (embed_dim=180, depths 6x6, heads 6, window 16) bf16 forward, "
                                   "batch 16 of 64x64 LR patches per GPU (BASELINE configs[1])",
                       "global_batch": BATCH * world, "lr_size": LR, "parallelism": f"image-batch shard x{world}",
                       "l2": "256 MiB buffer written between timed steps (L2 flush)",
                       "cuda_graph": bool(eng.use_cuda_graph), "tensor_cores": tc_on,
                       "weights": "random init, seed 0"},
            "clocks": clk,
            "e2e": {"value": e2e_val, "unit": "MP/s", "h2d_bytes_per_step": x_host.numel() * 2,
                    "d2h_bytes_per_step": y_host.numel() * 2},
            "gpu_launches": launches * K,
            "roofline": roof,
            "whole_forward": {"bound": "tensor", "achieved": whole, "peak": peaks["tf"], "unit": "TFLOP/s",
                              "frac": whole / peaks["tf"], "gflop_per_step": MFLOP_PER_LR_PX * lr_px / 1e3},
            "per_kernel": per_kernel, "per_op": per_op, "profiled_eager_step_ms": prof_ms,
            "tcgen05_fallbacks": eng.last_fallback_count,
        }
        if tiled is not None:
            line["tiled"] = tiled
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-tiled", action="store_true", help="skip the configs[3] tiled HAT-L measurement")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
