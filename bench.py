#!/usr/bin/env python
"""bench.py -- forward+backward transient rendering benchmark (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg2] [--mode analytic] [--impl reference]

A "step" = one fitting step of the hot path over one synthetic scene: per-Gaussian preprocess + shell binning +
batched forward H[scan_y,scan_x,T] + MSE gradient + backward + (N>1: one NCCL all-reduce of the packed
gradients) + parameter-gradient chain.  N=1 workload = BASELINE.json configs[1] (50k Gaussians, 64x64 scan,
512 bins, analytic integration).  N>1: every rank renders its own 64-row slab of a (64N)x64 scan grid (weak scaling).
One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOPS_PER_EVAL = {  # algorithmic FP32 flop per (Gaussian, ray, bin) sample-eval; MUFU counted separately
    ("numerical", "fwd"): 6, ("numerical", "bwd"): 14, ("analytic", "fwd"): 28, ("analytic", "bwd"): 44}
MUFU_PER_EVAL = {("numerical", "fwd"): 1, ("numerical", "bwd"): 1, ("analytic", "fwd"): 1, ("analytic", "bwd"): 2}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="cfg2")
    ap.add_argument("--mode", default="analytic", choices=["analytic", "numerical"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default): every rank renders its own Sy-row slab of a (Sy*N) x Sx grid; strong: the config's grid is split over the ranks")
    ap.add_argument("--variant", default="auto", choices=["auto", "wide", "narrow"], help="scan-tile kernel per entry (TransientRenderer)")
    ap.add_argument("--narrow-rays", type=float, default=2.5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-points", type=int, default=0, help="scan points in the CPU baseline sample (0 = auto)")
    return ap.parse_args()


class ClockSampler:
    """SM clock + throttle reasons sampled with NVML during the timed region."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.05)

    def __enter__(self):
        if self.nv:
            self.t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.nv:
            self.t.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def scene_for(cfg, n_gpus, strong=False):
    from nlos_gaussian_renderer_b200 import scenes
    N, Sy, Sx, T, A = scenes.CONFIGS[cfg]
    seed = 1000 + list(scenes.CONFIGS).index(cfg) + 1
    return (N, Sy if strong else Sy * n_gpus, Sx, T, A), seed



def ncu_traffic(cfg, mode):
    """dram__bytes_read.sum + dram__bytes_write.sum of the backward kernel, per launch, from the latest committed `ncu --set full` capture of
    the same workload (profiles/r01_<x>_ncu_raw_<cfg>_<mode>.txt, the last in name order); None when there is no capture for this workload."""
    import glob
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", f"r[0-9][0-9]_?_ncu_raw_{cfg}_{mode}.txt")))
    if not found:
        return None
    path = found[-1]
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = 0.0
    for ln in open(path):
        if ln.startswith("dram__bytes_read.sum") or ln.startswith("dram__bytes_write.sum"):
            u = ln.split("[")[1].split("]")[0]
            tot += float(ln.split("|")[2]) * unit.get(u, 1.0)          # column 2 = transient_kernel<mode, bwd>
    return tot or None


# ----------------------------------------------------------------------------------------------------------------
def cpu_baseline(cfg, mode, n_points, steps=1, warmup=0):
    """Times the reference's own CUDA-kernel source compiled for the host cores (oracle/_ref, kind "reference") when that
    library exists, else the C port (oracle/, kind "port"): forward only -- the reference backward returns zeros
    (gaussian_model/cuda_autograd.py:139-172) -- one render_rays call per scan point as main.py:198-203 does.
    Unit: Gaussian-bin evals/s with the bins counted like the GPU arm (W_b of the sampled scan points)."""
    import numpy as np
    import torch
    import oracle
    import oracle.ref as oref
    from nlos_gaussian_renderer_b200 import scenes
    from nlos_gaussian_renderer_b200.geometry import ScanGeometry
    shape, seed = scene_for(cfg, 1)
    N, Sy, Sx, T, A = shape
    geom = ScanGeometry.wall_grid(Sy, Sx, A=A)
    ta = scenes.time_axis(T)
    params = [p.numpy() for p in scenes.make_gaussians(N, seed)]
    use_ref = oref.available()
    cores = oref.num_threads() if use_ref else oracle.num_threads()
    if n_points <= 0:
        n_points = max(2, min(16, int(2.0e9 / (N * A * A))))
    idx = np.linspace(0, geom.S - 1, n_points).astype(int)
    t = ta["t_samples"].numpy()
    f32 = lambda x: x.numpy().astype(np.float32)
    so, st, ct, sp, cp = [f32(x) for x in (geom.scan_o, geom.sin_th, geom.cos_th, geom.sin_ph, geom.cos_ph)]

    def one_pass():
        for p in idx:
            d = np.stack([(st[p][:, None] * cp[p][None, :]).reshape(-1), (st[p][:, None] * sp[p][None, :]).reshape(-1),
                          np.repeat(ct[p], A)], 1).astype(np.float32)
            o = np.repeat(so[p][None], A * A, 0)
            if use_ref:
                oref.render_rays(o, d, t, *params, so[p], 0, ta["c"], ta["deltaT"], 1.0, False)
            else:
                oracle.render_rays(o, d, t, *params, so[p], 0, ta["c"], ta["deltaT"], 1.0, False, cap=256, exp_mode=0)

    # work of the sample, counted with the same definition as the GPU arm (uncapped W_b); bounded: 2 scan points
    sub = idx[:2]
    _, cnt = oracle.transient_batch(so[sub], st[sub], ct[sub], sp[sub], cp[sub], t, f32(geom.hscale)[sub], *params, sh_degree=0,
                                    c=ta["c"], deltaT=ta["deltaT"], count=True)
    wb_per_point = cnt["W_b"] / len(sub)
    for _ in range(warmup):
        one_pass()
    t0 = time.perf_counter()
    for _ in range(steps):
        one_pass()
    el = (time.perf_counter() - t0) / steps
    value = wb_per_point * n_points / el
    return {"value": value, "unit": "Gaussian-bin evals/s", "cores": cores, "kind": "reference" if use_ref else "port",
            "sample": f"forward only (reference has no backward), {n_points} of {geom.S} scan points of {cfg}, one render_rays "
                      f"call per scan point, 256-per-ray cap as in the reference, {el / n_points * 1e3:.1f} ms/scan point",
            "ms_per_scan_point": el / n_points * 1e3, "seconds": el}


def run_reference(args):
    """--impl reference: the reference's own CPU-runnable implementation of the path (its CUDA kernel source compiled for
    the host, oracle/_ref; else the C port) on all host threads.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    steps, warm = max(args.steps, 1), args.warmup
    # bounded: each step is a sample of scan points sized so the whole run ends within a few minutes
    b = cpu_baseline(args.config, args.mode, args.cpu_points or 4, steps=steps, warmup=min(warm, 1))
    (N, Sy, Sx, T, A), _ = scene_for(args.config, max(args.gpus, 1), args.scaling == "strong")
    workload = (f"{args.config}: {N} Gaussians, {Sy}x{Sx} confocal scan ({Sy // max(args.gpus, 1)} rows per GPU), {T} bins, A={A}, "
                f"{args.mode} integration, forward+backward fitting step")
    line = {"impl": "reference", "metric": "Gaussian-bin evals/s (forward+backward transient)", "value": b["value"],
            "unit": "Gaussian-bin evals/s", "n_gpus": args.gpus, "steps": steps, "warmup": min(warm, 1),
            "ms_per_step": b["seconds"] * 1e3, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": workload, "sample": b["sample"],
                                             "note": "the reference's backward is a zero stub (cuda_autograd.py:139-172) and its numerical kernel is its only "
                                                     "per-bin renderer: the step timed here is its forward; bounded sample of scan points"},
            "cpu_baseline": {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": b["value"], "unit": "Gaussian-bin evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def run_b200(args):
    import torch
    import torch.distributed as dist
    from nlos_gaussian_renderer_b200 import scenes, parallel
    from nlos_gaussian_renderer_b200.geometry import ScanGeometry
    from nlos_gaussian_renderer_b200.transient import TransientRenderer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    shape, seed = scene_for(args.config, world, args.scaling == "strong")
    N, Sy, Sx, T, A = shape
    ta = scenes.time_axis(T)
    full = ScanGeometry.wall_grid(Sy, Sx, A=A)
    y0, y1 = parallel.slab_rows(Sy, full.tile_y, rank, world)
    g0 = full.rows(y0, y1)
    geom = ScanGeometry(g0.scan_o.reshape(y1 - y0, Sx, 3), g0.theta_range.reshape(y1 - y0, Sx, 2), g0.phi_range.reshape(y1 - y0, Sx, 2),
                        A, vp_y=full.vp_y, device=dev)
    host_params = [p.pin_memory() for p in scenes.make_gaussians(N, seed)]
    params = [p.to(dev) for p in host_params]
    rnd = TransientRenderer(geom, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=args.mode, variant=args.variant, narrow_rays=args.narrow_rays)
    target = rnd.forward(*[p.to(dev) for p in scenes.make_gaussians(N, seed + 1)]).clone()      # the "measured" transient
    fit = parallel.ShardedFitter(rnd, target, Sy * Sx * T)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ev = lambda: torch.cuda.Event(enable_timing=True)
    kt = {"fwd": [], "bwd": []}

    def timed_step(record):
        flush.zero_()                                                    # L2 flush between iterations (256 MB > 126 MB L2)
        rnd.prepare(*params)
        e = [ev() for _ in range(4)]
        e[0].record(); Ht = rnd.forward_tiled(); e[1].record()
        loss, gH = rnd.mse_loss_grad_tiled(Ht, fit.target_tiled, 1.0 / fit.numel)      # fused loss + dL/dH (csrc/loss.cu)
        e[2].record()
        rnd.render_backward(gH)
        e[3].record()
        parallel.all_reduce_packed(rnd.dacc)
        outs = rnd.chain()
        if record:
            kt["fwd"].append((e[0], e[1]))
            kt["bwd"].append((e[2], e[3]))
        return loss, outs

    for _ in range(max(args.warmup, 3)):
        timed_step(False)
    work = rnd.count_work()
    sync()
    with ClockSampler(local) as clk:
        a, b = ev(), ev()
        a.record()
        for _ in range(args.steps):
            loss, grads = timed_step(True)
        b.record()
        sync()
    ms_total = a.elapsed_time(b)
    # ---- end-to-end through the public API with HOST buffers
    hg = [torch.empty_like(p).pin_memory() for p in host_params]
    hl = torch.empty((), dtype=torch.float32).pin_memory()
    sync()
    c0, c1 = ev(), ev()
    c0.record()
    for _ in range(args.steps):
        dp = [p.to(dev, non_blocking=True) for p in host_params]
        loss, grads = fit.step(*dp)
        for h, g_ in zip(hg, grads):
            h.copy_(g_.reshape(h.shape), non_blocking=True)
        hl.copy_(loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()
    c1.record()
    sync()
    ms_e2e = c0.elapsed_time(c1)
    t = torch.tensor([ms_total, ms_e2e], dtype=torch.float64, device=dev)
    w = torch.tensor([work["W_b"], work["W_s"], work["evals"], work["hits"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
    ms_total, ms_e2e = t.tolist()
    Wb, Ws, Wev, hits = w.tolist()
    if rank == 0:
        ms_step = ms_total / args.steps
        fwd_ms = sum(x.elapsed_time(y) for x, y in kt["fwd"]) / len(kt["fwd"])
        bwd_ms = sum(x.elapsed_time(y) for x, y in kt["bwd"]) / len(kt["bwd"])
        clocks = clk.summary()
        f_mhz = clocks["sm_mhz"] or 1965
        peak_fp32 = 148 * 128 * 2 * f_mhz * 1e6 / 1e12                    # TFLOP/s at the SM clock seen during the run
        peak_sfu = 148 * 16 * f_mhz * 1e6                                  # MUFU ops/s
        ws_rank = work["W_s"]                                               # rank 0's kernel

        def roof(kind, ms):
            """achieved = W_s / t against min(FP32 ceiling, SFU ceiling) for the algorithmic op mix of one sample-eval,
            expressed in TFLOP/s: peak = min(FP32 peak, SFU rate x flop-per-MUFU of the mix)."""
            fl, mu = FLOPS_PER_EVAL[(args.mode, kind)], MUFU_PER_EVAL[(args.mode, kind)]
            ach = ws_rank * fl / (ms * 1e-3) / 1e12
            sfu_equiv = peak_sfu * fl / mu / 1e12
            peak = min(peak_fp32, sfu_equiv)
            return {"launch_ms": ms, "achieved": ach, "peak": peak, "frac": ach / peak, "binding_pipe": "fp32" if peak_fp32 <= sfu_equiv else "sfu",
                    "fp32_frac": ach / peak_fp32, "sfu_frac": ws_rank * mu / (ms * 1e-3) / peak_sfu, "flop_per_sample_eval": fl,
                    "mufu_per_sample_eval": mu, "sample_evals_per_s": ws_rank / (ms * 1e-3)}

        rb, rf = roof("bwd", bwd_ms), roof("fwd", fwd_ms)
        traffic = ncu_traffic(args.config, args.mode) if world == 1 else None
        hbm_view = None
        try:      # the HBM view of the same launch against the driver-measured copy bandwidth: shows how far from HBM-bound the kernel is
            hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
            hbm_src = "MEASURED_PEAKS.json hbm_gbs"
        except Exception:
            hbm_peak, hbm_src = 6550.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"
        if traffic:
            gbs = traffic / (bwd_ms * 1e-3) / 1e9
            hbm_view = {"achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "peak_source": hbm_src,
                        "bytes_per_launch": traffic}
        n_bytes = N * 12 * 4
        line = {
            "metric": "Gaussian-bin evals/s (forward+backward transient)", "value": Wb / (ms_step * 1e-3), "unit": "Gaussian-bin evals/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.config}: {N} Gaussians, {Sy}x{Sx} confocal scan ({Sy // world} rows per GPU), {T} bins, A={A}, "
                                   f"{args.mode} integration, forward+backward fitting step", "parallelism": f"scan-slab x{world}",
                       "l2": "256 MB memset between steps (inside the timed region)", "mode": args.mode,
                       "W_b": Wb, "W_s": Ws, "sample_evals_evaluated": Wev, "hit_ray_pairs": hits},
            "fwd_bwd_ms_per_transient": ms_step, "fwd_kernel_ms": fwd_ms, "bwd_kernel_ms": bwd_ms,
            "sample_evals_per_s": Ws / (ms_step * 1e-3),
            "clocks": clocks,
            "e2e": {"value": Wb / (ms_e2e / args.steps * 1e-3), "unit": "Gaussian-bin evals/s", "h2d_bytes_per_step": n_bytes,
                    "d2h_bytes_per_step": n_bytes + 4, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": 9 * args.steps,
            "roofline": {"bound": "compute (" + rb["binding_pipe"] + " pipe)", "kernel": f"transient_kernel<{args.mode},bwd>", "achieved": rb["achieved"],
                         "peak": rb["peak"], "unit": "TFLOP/s", "frac": rb["frac"],
                         "traffic": traffic, "hbm": hbm_view,
                         "peak_source": f"min(148 SM x 128 FP32 lanes x 2 flop, 148 SM x 16 MUFU/clk x flop-per-MUFU of the algorithmic mix) x {f_mhz} MHz "
                                        "(SM clock sampled during the run; SURVEY.md 8d); MEASURED_PEAKS.json holds only HBM and bf16 tensor peaks, "
                                        "neither binds this kernel (DRAM traffic ~75 MB per launch: see the hbm view)",
                         "sample_evals_per_launch": ws_rank, "bwd": rb, "fwd": rf},
            "loss": float(loss),
        }
        if world == 1 and not args.no_cpu_baseline:
            b = cpu_baseline(args.config, args.mode, args.cpu_points)
            line["cpu_baseline"] = {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
