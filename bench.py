#!/usr/bin/env python
"""bench.py -- forward+backward transient rendering benchmark (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg4] [--mode analytic] [--scaling strong] [--impl reference]

A "step" = one fitting step of the hot path over one synthetic scene: per-Gaussian preprocess + shell binning +
batched forward H[scan_y,scan_x,T] + fused MSE loss / dL/dH + backward + (N>1: ONE NCCL all-reduce of the packed
gradients) + parameter-gradient chain.

Workload: BASELINE.json configs[3] (the fitting loop the metric is quoted on at 1/2/4/8 B200: 500k Gaussians -> 128x128x1024,
the largest configuration that is specified on ONE GPU), analytic integration.  N>1 is STRONG scaling by default: the same
scene, its scan rows dealt round-robin (tile rows) over the ranks.  `--scaling weak` gives every rank its own copy of the grid.
Rank 0 prints ONE JSON line.  Extra blocks in the same line (N=1): `north_star` (1M Gaussians -> 128x128x1024, both
integration modes, fwd / bwd ms and roofline fraction), `numerical_forward_only` (the mode the reference arm runs),
`cpu_baseline` (the reference's own kernel source on the host cores) and `cpu_dense_torch` (its dense PyTorch path).
N>1 adds `grad_check`: the all-reduced packed gradient against rank 0's own single-GPU render of the full grid.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# algorithmic FP32 flop / MUFU per (Gaussian, ray, bin) sample-eval (DESIGN.md 5 derives the analytic figures)
FLOPS_PER_EVAL = {("numerical", "fwd"): 6, ("numerical", "bwd"): 14, ("analytic", "fwd"): 28, ("analytic", "bwd"): 44}
MUFU_PER_EVAL = {("numerical", "fwd"): 1, ("numerical", "bwd"): 1, ("analytic", "fwd"): 1, ("analytic", "bwd"): 2}
METRIC = "Gaussian-bin evals/s (forward+backward transient)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="cfg4")
    ap.add_argument("--mode", default="analytic", choices=["analytic", "numerical"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="strong (default): the config's scan grid is dealt over the ranks; weak: every rank renders its own copy of the grid rows")
    ap.add_argument("--variant", default="auto", choices=["auto", "wide", "narrow"], help="scan-tile kernel per entry (TransientRenderer)")
    ap.add_argument("--narrow-rays", type=float, default=2.5)
    ap.add_argument("--contiguous", action="store_true", help="contiguous row slabs instead of round-robin tile rows")
    ap.add_argument("--occlusion", action="store_true", help="add the `occlusion` block: the batched occlusion renderer (csrc/occluded.cu) on BASELINE configs[1]")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the north_star / numerical_forward_only / grad_check blocks")
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


def scene_for(cfg, n_gpus, strong=True):
    from nlos_gaussian_renderer_b200 import scenes
    N, Sy, Sx, T, A = scenes.CONFIGS[cfg]
    seed = 1000 + list(scenes.CONFIGS).index(cfg) + 1
    return (N, Sy if strong else Sy * n_gpus, Sx, T, A), seed


def workload_name(cfg, shape, world, mode, scaling):
    N, Sy, Sx, T, A = shape
    return (f"{cfg}: {N} Gaussians, {Sy}x{Sx} confocal scan, {T} bins, A={A}, {mode} integration, forward+backward fitting step"
            + (f"; {scaling} scaling over {world} GPUs" if world > 1 else ""))


def ncu_traffic(cfg, mode):
    """(bytes, tag): dram__bytes_read.sum + dram__bytes_write.sum of the dominant BACKWARD kernel launch (the longest one), from the newest
    COMMITTED `ncu --set full` capture of this workload (profiles/rNN_x_ncu_raw_<cfg>_<mode>.txt, written by tools/ncu_summary.py raw).
    It is not a measurement of this run: the tag says which capture."""
    import glob
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", f"r[0-9][0-9]_?_ncu_raw_{cfg}_{mode}.txt")))
    if not found:
        return None, None
    path = found[-1]
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    names, dur, rd, wr = [], [], [], []
    num = lambda c: float(c) if c.strip().replace(".", "", 1).replace("e+", "", 1).replace("e-", "", 1).isdigit() else float("nan")
    for ln in open(path):
        cols = [c.strip() for c in ln.rstrip("\n").split("|")]
        if ln.startswith("metric |"):
            names = cols[1:]
        elif ln.startswith("gpu__time_duration.sum"):
            dur = [num(c) for c in cols[1:]]
        elif ln.startswith("dram__bytes_read.sum") or ln.startswith("dram__bytes_write.sum"):
            u = unit.get(ln.split("[")[1].split("]")[0], 1.0)
            vals = [num(c) * u for c in cols[1:]]
            if ln.startswith("dram__bytes_read"):
                rd = vals
            else:
                wr = vals
    tag = os.path.basename(path)
    for want_bwd in (True, False):                               # (a capture whose backward counters came back NaN: the dominant forward launch, labelled)
        best, tot = -1.0, None
        for i, n in enumerate(names):
            is_bwd = ", 1, " in n or ", 1>" in n                 # transient[_narrow]_kernel<MODE, BWD = 1, SH>
            if is_bwd == want_bwd and i < len(dur) and i < len(rd) and i < len(wr) and dur[i] == dur[i] and rd[i] == rd[i] and dur[i] > best:
                best, tot = dur[i], rd[i] + (wr[i] if wr[i] == wr[i] else 0.0)
        if tot is not None:
            return tot, tag + ("" if want_bwd else ": dominant FORWARD launch (the backward's dram counters are NaN in this capture)")
    return None, tag


# ----------------------------------------------------------------------------------------------------------------
def _host_scene(cfg):
    import numpy as np
    from nlos_gaussian_renderer_b200 import scenes
    from nlos_gaussian_renderer_b200.geometry import ScanGeometry
    shape, seed = scene_for(cfg, 1)
    N, Sy, Sx, T, A = shape
    geom = ScanGeometry.wall_grid(Sy, Sx, A=A)
    ta = scenes.time_axis(T)
    params = [p.numpy() for p in scenes.make_gaussians(N, seed)]
    f32 = lambda x: x.numpy().astype(np.float32)
    tabs = [f32(x) for x in (geom.scan_o, geom.sin_th, geom.cos_th, geom.sin_ph, geom.cos_ph)]
    return shape, geom, ta, params, tabs


def cpu_baseline(cfg, n_points, steps=1, warmup=0, budget_s=20.0):
    """Times the reference's own CUDA-kernel source compiled for the host cores (oracle/_ref, kind "reference") when that
    library exists, else the C port (oracle/, kind "port"), on ALL host threads: forward only -- the reference backward returns zeros
    (gaussian_model/cuda_autograd.py:139-172) -- one render_rays call per scan point as main.py:198-203 does.
    Unit: Gaussian-bin evals/s with the bins counted like the GPU arm (W_b of the sampled scan points).
    n_points <= 0: one scan point is timed first and the sample is sized so that (steps + warmup) passes fit budget_s."""
    import numpy as np
    import oracle
    import oracle.ref as oref
    shape, geom, ta, params, (so, st, ct, sp, cp) = _host_scene(cfg)
    N, Sy, Sx, T, A = shape
    use_ref = oref.available()
    ncpu = os.cpu_count() or 1
    (oref if use_ref else oracle).set_num_threads(ncpu)          # (torchrun exports OMP_NUM_THREADS=1: ask for every core explicitly)
    oracle.set_num_threads(ncpu)
    cores = oref.num_threads() if use_ref else oracle.num_threads()
    t = ta["t_samples"].numpy()

    def one_point(p):
        d = np.stack([(st[p][:, None] * cp[p][None, :]).reshape(-1), (st[p][:, None] * sp[p][None, :]).reshape(-1),
                      np.repeat(ct[p], A)], 1).astype(np.float32)
        o = np.repeat(so[p][None], A * A, 0)
        if use_ref:
            oref.render_rays(o, d, t, *params, so[p], 0, ta["c"], ta["deltaT"], 1.0, False)
        else:
            oracle.render_rays(o, d, t, *params, so[p], 0, ta["c"], ta["deltaT"], 1.0, False, cap=256, exp_mode=0)

    if n_points <= 0:
        t0 = time.perf_counter()
        one_point(geom.S // 2)
        t1 = time.perf_counter() - t0
        n_points = int(max(1, min(16, budget_s / (max(steps + warmup, 1) * t1))))
    idx = np.linspace(0, geom.S - 1, n_points + 2).astype(int)[1:-1] if n_points < geom.S else np.arange(geom.S)
    # work of the sample, counted with the same definition as the GPU arm (uncapped W_b); bounded: 2 scan points
    sub = idx[:2]
    _, cnt = oracle.transient_batch(so[sub], st[sub], ct[sub], sp[sub], cp[sub], t, geom.hscale.numpy().astype(np.float32)[sub], *params,
                                    sh_degree=0, c=ta["c"], deltaT=ta["deltaT"], count=True)
    wb_per_point = cnt["W_b"] / len(sub)
    for _ in range(warmup):
        for p in idx:
            one_point(p)
    t0 = time.perf_counter()
    for _ in range(steps):
        for p in idx:
            one_point(p)
    el = (time.perf_counter() - t0) / steps
    value = wb_per_point * len(idx) / el
    return {"value": value, "unit": "Gaussian-bin evals/s", "cores": cores, "kind": "reference" if use_ref else "port",
            "sample": f"forward only, numerical integration (the reference has no backward and no per-bin analytic mode), {len(idx)} of {geom.S} scan "
                      f"points of {cfg}, one render_rays call per scan point, 256-per-ray cap as in the reference, {el / len(idx) * 1e3:.1f} ms/scan point "
                      f"on {cores} threads",
            "ms_per_scan_point": el / len(idx) * 1e3, "seconds": el}


def cpu_dense_torch(cfg, budget_s=15.0):
    """The reference's DENSE PyTorch path (nlos_helpers.py:206-232 -> gaussian_model.py:253-294,346-364; the path that trains with
    use_cuda_renderer=False) through its port oracle/dense_torch.py (pinned against the reference's own output in tests/test_ref_python.py),
    on all host threads: 8 scan points of the config's lattice x a SUBSET of the Gaussians sized for budget_s, forward and
    forward+backward; the dense cost is exactly linear in the number of Gaussians (a sum over Gaussians), so the per-scan-point time is
    extrapolated to the config's N and stated as such.  Unit: dense Gaussian-sample evals/s (N * A^2 * T per scan point)."""
    import torch
    from oracle import dense_torch
    shape, geom, ta, params, _ = _host_scene(cfg)
    N, Sy, Sx, T, A = shape
    ncpu = os.cpu_count() or 1
    torch.set_num_threads(ncpu)
    pts_idx = torch.linspace(0, geom.S - 1, 10).long()[1:-1]                   # 8 scan points
    r = ta["t_samples"]
    rate_guess = 4.0e7 / 8 * ncpu                                               # SURVEY App. B probe: 4e7 dense evals/s forward on 8 cores
    G = int(max(8, min(N, budget_s * rate_guess / 3.5 / (8 * A * A * T))))      # fwd + (fwd+bwd ~ 2.5 x fwd)
    P = [torch.from_numpy(p[:G]).clone() for p in params]
    # (dense semantics: scale = exp(exp(_scaling)); the timing does not depend on the values)
    tf = tb = 0.0
    for p in pts_idx.tolist():
        th, ph = geom.theta[p], geom.phi[p]
        o = geom.scan_o[p]
        pts = torch.stack([r[:, None, None] * torch.sin(th)[None, :, None] * torch.cos(ph)[None, None, :],
                           r[:, None, None] * torch.sin(th)[None, :, None] * torch.sin(ph)[None, None, :],
                           (r[:, None, None] * torch.cos(th)[None, :, None]).expand(T, A, A)], -1) + o
        dth = float(geom.theta_range[p, 1] - geom.theta_range[p, 0]) / A
        dph = float(geom.phi_range[p, 1] - geom.phi_range[p, 0]) / A
        chunk = max(1, min(64, int(2.0e8 / (A * A * T))))
        t0 = time.perf_counter()
        with torch.no_grad():
            dense_torch.dense_histogram(pts, th, r, *P, dth, dph, geom.vp_y, chunk=chunk)
        tf += time.perf_counter() - t0
        Q = [q.clone().requires_grad_() for q in P]
        t0 = time.perf_counter()
        h = dense_torch.dense_histogram(pts, th, r, *Q, dth, dph, geom.vp_y, chunk=chunk)
        h.sum().backward()
        tb += time.perf_counter() - t0
    n_eval = G * A * A * T * len(pts_idx)
    return {"value": n_eval / tf, "unit": "dense Gaussian-sample evals/s (forward)", "fwd_bwd_value": n_eval / tb, "cores": ncpu, "kind": "port",
            "sample": f"oracle/dense_torch.py (port of gaussian_model.py:253-294,346-364 + nlos_helpers.py:206-232), {len(pts_idx)} scan points x {G} of {N} "
                      f"Gaussians x {A}x{A}x{T} lattice samples on {ncpu} torch threads; linear in N, extrapolated",
            "ms_per_scan_point_fwd_extrapolated": tf / len(pts_idx) * 1e3 * N / G,
            "ms_per_scan_point_fwd_bwd_extrapolated": tb / len(pts_idx) * 1e3 * N / G,
            "s_per_transient_fwd_bwd_extrapolated": tb / len(pts_idx) * N / G * geom.S}


def run_reference(args):
    """--impl reference: the reference's own CPU-runnable implementation of the path (its CUDA kernel source compiled for
    the host, oracle/_ref; else the C port) on ALL host threads.  Rank 0 only.  Each step is a bounded sample of the config's scan points,
    sized after timing one scan point so that the whole --steps K --warmup W run ends within ~2 minutes."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    steps, warm = max(args.steps, 1), max(args.warmup, 0)
    b = cpu_baseline(args.config, args.cpu_points, steps=steps, warmup=warm, budget_s=100.0)
    shape, _ = scene_for(args.config, max(args.gpus, 1), args.scaling == "strong")
    line = {"impl": "reference", "metric": METRIC, "value": b["value"],
            "unit": "Gaussian-bin evals/s", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": b["seconds"] * 1e3, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": workload_name(args.config, shape, max(args.gpus, 1), args.mode, args.scaling), "sample": b["sample"],
                                             "note": "the reference's backward is a zero stub (cuda_autograd.py:139-172) and its numerical kernel is its only "
                                                     "per-bin renderer: the step timed here is its numerical FORWARD on a bounded sample of scan points; the GPU arm's "
                                                     "`numerical_forward_only` block is the like-for-like leg"},
            "cpu_baseline": {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": b["value"], "unit": "Gaussian-bin evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------------------
def roofline_of(mode, kind, ws, ms, f_mhz):
    """achieved = W_s / t against min(FP32 ceiling, SFU ceiling) for the ALGORITHMIC op mix of one sample-eval,
    expressed in TFLOP/s: peak = min(FP32 peak, SFU rate x flop-per-MUFU of the mix)."""
    peak_fp32 = 148 * 128 * 2 * f_mhz * 1e6 / 1e12
    peak_sfu = 148 * 16 * f_mhz * 1e6
    fl, mu = FLOPS_PER_EVAL[(mode, kind)], MUFU_PER_EVAL[(mode, kind)]
    ach = ws * fl / (ms * 1e-3) / 1e12
    sfu_equiv = peak_sfu * fl / mu / 1e12
    peak = min(peak_fp32, sfu_equiv)
    return {"launch_ms": ms, "achieved": ach, "peak": peak, "frac": ach / peak, "binding_pipe": "fp32" if peak_fp32 <= sfu_equiv else "sfu",
            "fp32_frac": ach / peak_fp32, "sfu_frac": ws * mu / (ms * 1e-3) / peak_sfu, "flop_per_sample_eval": fl,
            "mufu_per_sample_eval": mu, "sample_evals_per_s": ws / (ms * 1e-3)}


def time_kernels(rnd, params, target_tiled, inv_numel, steps, warmup, flush):
    """-> (fwd ms, bwd ms) of the render kernels alone (CUDA events on the launching stream), averaged over `steps`."""
    import torch
    ev = lambda: torch.cuda.Event(enable_timing=True)
    tf = tb = 0.0
    for it in range(warmup + steps):
        flush.zero_()
        rnd.prepare(*params)
        e = [ev() for _ in range(4)]
        e[0].record(); Ht = rnd.forward_tiled(); e[1].record()
        _, gH = rnd.mse_loss_grad_tiled(Ht, target_tiled if target_tiled is not None else torch.zeros_like(Ht), inv_numel)
        e[2].record(); rnd.render_backward(gH); e[3].record()
        rnd.chain()
        torch.cuda.synchronize()
        if it >= warmup:
            tf += e[0].elapsed_time(e[1]); tb += e[2].elapsed_time(e[3])
    return tf / steps, tb / steps


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
    rows = parallel.rank_rows(Sy, full.tile_y, rank, world, interleave=not args.contiguous)
    sub = full.select_rows(rows)
    geom = ScanGeometry(sub.scan_o.reshape(len(rows), Sx, 3), sub.theta_range.reshape(len(rows), Sx, 2), sub.phi_range.reshape(len(rows), Sx, 2),
                        A, vp_y=full.vp_y, device=dev)
    host_params = [p.pin_memory() for p in scenes.make_gaussians(N, seed)]
    params = [p.to(dev) for p in host_params]
    mk = lambda g, mode: TransientRenderer(g, T, ta["t0"], ta["dt"], ta["c"] * ta["deltaT"], mode=mode, variant=args.variant, narrow_rays=args.narrow_rays)
    rnd = mk(geom, args.mode)
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

    W = max(args.warmup, 3)
    for _ in range(W):
        timed_step(False)
    rnd.check_lists()
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
    rnd.check_lists()
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
    fwd_ms = sum(x.elapsed_time(y) for x, y in kt["fwd"]) / len(kt["fwd"])
    bwd_ms = sum(x.elapsed_time(y) for x, y in kt["bwd"]) / len(kt["bwd"])
    t = torch.tensor([ms_total, ms_e2e, fwd_ms, bwd_ms], dtype=torch.float64, device=dev)
    w = torch.tensor([work["W_b"], work["W_s"], work["evals"], work["hits"]], dtype=torch.float64, device=dev)
    tmin = t.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
    ms_total, ms_e2e, fwd_max, bwd_max = t.tolist()
    Wb, Ws, Wev, hits = w.tolist()

    # ---- N>1, untimed: the all-reduced packed gradient == rank 0's own single-GPU render of the full grid
    grad_check = None
    if world > 1 and not args.no_extra:
        rnd.prepare(*params)
        Ht = rnd.forward_tiled()
        _, gH = rnd.mse_loss_grad_tiled(Ht, fit.target_tiled, 1.0 / fit.numel)
        rnd.render_backward(gH)
        parallel.all_reduce_packed(rnd.dacc)
        reduced = rnd.dacc.clone()
        tgt_full = parallel.gather_transient(target, Sy, full.tile_y, interleave=not args.contiguous)
        if rank == 0:
            gfull = ScanGeometry(full.scan_o.reshape(Sy, Sx, 3), full.theta_range.reshape(Sy, Sx, 2), full.phi_range.reshape(Sy, Sx, 2), A,
                                 vp_y=full.vp_y, device=dev)
            r1 = mk(gfull, args.mode)
            r1.prepare(*params)
            H1 = r1.forward_tiled()
            _, g1 = r1.mse_loss_grad_tiled(H1, gfull.tile_grad(tgt_full).contiguous(), 1.0 / fit.numel)
            r1.render_backward(g1)
            err = float((r1.dacc - reduced).abs().max() / r1.dacc.abs().max().clamp_min(1e-30))
            grad_check = {"max_abs_err_over_max": err, "ok": err <= 1e-5,
                          "what": "all-reduced packed gradient [N,16] of the sharded step vs rank 0 rendering the whole grid alone (untimed)"}
            del r1, H1, g1
        sync()

    line = None
    if rank == 0:
        ms_step = ms_total / args.steps
        clocks = clk.summary()
        f_mhz = clocks["sm_mhz"] or 1965
        ws_rank = work["W_s"]                                               # rank 0's kernel
        rb, rf = roofline_of(args.mode, "bwd", ws_rank, bwd_ms, f_mhz), roofline_of(args.mode, "fwd", ws_rank, fwd_ms, f_mhz)
        traffic, ttag = ncu_traffic(args.config, args.mode) if world == 1 else (None, None)
        try:      # the HBM view of the same launch against the driver-measured copy bandwidth: shows how far from HBM-bound the kernel is
            hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
            hbm_src = "MEASURED_PEAKS.json hbm_gbs"
        except Exception:
            hbm_peak, hbm_src = 6650.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"
        alg_bytes = 48 * N + 4 * geom.S * T + 4 * 12 * N                   # parameters + dL/dH + packed gradients
        hbm_view = {"algorithmic_bytes_per_launch": alg_bytes, "algorithmic_gbs": alg_bytes / (bwd_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "peak_source": hbm_src}
        if traffic:
            t_ms = fwd_ms if "FORWARD" in (ttag or "") else bwd_ms       # (the pass the captured launch belongs to)
            hbm_view.update({"achieved": traffic / (t_ms * 1e-3) / 1e9, "frac": traffic / (t_ms * 1e-3) / 1e9 / hbm_peak, "bytes_per_launch": traffic})
        n_bytes = N * 12 * 4
        n_cls = len(rnd._launch_groups())        # one launch per kernel variant (its entry classes are segments of the launch)
        n_flt = 1 if rnd.one_pass else 0
        line = {
            "metric": METRIC, "value": Wb / (ms_step * 1e-3), "unit": "Gaussian-bin evals/s",
            "n_gpus": world, "steps": args.steps, "warmup": W, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.config, shape, world, args.mode, args.scaling),
                       "parallelism": f"scan rows x{world}" + ("" if world == 1 else (", contiguous slabs" if args.contiguous else ", tile rows round-robin")),
                       "l2": "256 MB memset between steps (inside the timed region)", "mode": args.mode, "kernel_variant": args.variant,
                       "W_b": Wb, "W_s": Ws, "sample_evals_evaluated": Wev, "hit_ray_pairs": hits},
            "fwd_bwd_ms_per_transient": ms_step, "fwd_kernel_ms": fwd_ms, "bwd_kernel_ms": bwd_ms,
            "kernel_ms_over_ranks": {"fwd_max": fwd_max, "fwd_min": tmin[2].item(), "bwd_max": bwd_max, "bwd_min": tmin[3].item()},
            "sample_evals_per_s": Ws / (ms_step * 1e-3),
            "clocks": clocks,
            "e2e": {"value": Wb / (ms_e2e / args.steps * 1e-3), "unit": "Gaussian-bin evals/s", "h2d_bytes_per_step": n_bytes,
                    "d2h_bytes_per_step": n_bytes + 4, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": (1 + 4 + n_cls + n_flt + 1 + n_flt + n_cls + 1) * args.steps,
            "gpu_launches_per_step": {"preprocess": 1, "binning": 4, "transient_forward": n_cls, "analytic_filter": 2 * n_flt, "loss": 1, "transient_backward": n_cls,
                                      "chain": 1, "entry_classes": rnd.n_classes},
            "roofline": {"bound": "compute (" + rb["binding_pipe"] + " pipe)", "kernel": f"transient kernels <{args.mode},bwd> ({args.variant})",
                         "achieved": rb["achieved"], "peak": rb["peak"], "unit": "TFLOP/s", "frac": rb["frac"],
                         "frac_kind": "algorithmic-equivalent (SURVEY 8d: W_s x the algorithmic flop of one sample-eval / launch time / pipe peak); "
                                      "the kernels execute a cheaper recurrence, so this is NOT pipe utilisation -- ncu's executed view is in profiles/",
                         "traffic": traffic, "traffic_source": (f"committed capture profiles/{ttag} (not measured in this run)" if ttag else None),
                         "hbm": hbm_view,
                         "peak_source": f"min(148 SM x 128 FP32 lanes x 2 flop, 148 SM x 16 MUFU/clk x flop-per-MUFU of the algorithmic mix) x {f_mhz} MHz "
                                        "(SM clock sampled during the run; SURVEY.md 8d); MEASURED_PEAKS.json holds only HBM and bf16 tensor peaks, "
                                        "neither binds this kernel (see the hbm view)",
                         "sample_evals_per_launch": ws_rank, "bwd": rb, "fwd": rf},
            "loss": float(loss),
        }
        if grad_check is not None:
            line["grad_check"] = grad_check
    if world > 1:
        dist.destroy_process_group()
    if rank != 0:
        return

    if world == 1 and not args.no_extra:
        del fit, target
        f_mhz = line["clocks"]["sm_mhz"] or 1965
        # ---- the mode the reference arm runs: numerical integration, forward only (like-for-like leg of the CPU ratio)
        if args.mode == "numerical":
            nf_ms, nwb = fwd_ms, Wb
        else:
            rn = mk(geom, "numerical")
            nf_ms, _ = time_kernels(rn, params, None, 1.0, 3, 2, flush)
            nwb = rn.count_work()["W_b"]
            del rn
        line["numerical_forward_only"] = {"value": nwb / (nf_ms * 1e-3), "unit": "Gaussian-bin evals/s", "fwd_kernel_ms": nf_ms,
                                          "what": "numerical integration, forward kernel only, same scene: the configuration the CPU reference arm times"}
        del rnd
        torch.cuda.empty_cache()
        # ---- the north-star target: 1M Gaussians -> 128x128x1024 on one GPU, both integration modes
        ns = {}
        Nn, Syn, Sxn, Tn, An = scenes.CONFIGS["north_star"]
        gn = ScanGeometry.wall_grid(Syn, Sxn, A=An)
        gn = ScanGeometry(gn.scan_o.reshape(Syn, Sxn, 3), gn.theta_range.reshape(Syn, Sxn, 2), gn.phi_range.reshape(Syn, Sxn, 2), An, vp_y=gn.vp_y, device=dev)
        tan = scenes.time_axis(Tn)
        pn = [p.to(dev) for p in scenes.make_gaussians(Nn, 1000 + list(scenes.CONFIGS).index("north_star") + 1)]
        for mode in ("numerical", "analytic"):
            rn = TransientRenderer(gn, Tn, tan["t0"], tan["dt"], tan["c"] * tan["deltaT"], mode=mode, variant=args.variant, narrow_rays=args.narrow_rays)
            f_ms, b_ms = time_kernels(rn, pn, None, 1.0 / (Syn * Sxn * Tn), 3, 2, flush)
            wk = rn.count_work()
            ns[mode] = {"fwd_kernel_ms": f_ms, "bwd_kernel_ms": b_ms, "W_s": wk["W_s"], "W_b": wk["W_b"],
                        "gaussian_bin_evals_per_s": wk["W_b"] / ((f_ms + b_ms) * 1e-3),
                        "frac_fwd": roofline_of(mode, "fwd", wk["W_s"], f_ms, f_mhz)["frac"], "frac_bwd": roofline_of(mode, "bwd", wk["W_s"], b_ms, f_mhz)["frac"]}
            del rn
            torch.cuda.empty_cache()
        line["north_star"] = {"workload": f"{Nn} Gaussians, {Syn}x{Sxn} confocal scan, {Tn} bins, A={An}: forward + backward kernels on ONE B200 (3 steps after 2 warm-ups)",
                              "frac_kind": "algorithmic-equivalent, as roofline.frac", **ns}
    if world == 1 and args.occlusion:
        # ---- the reference's per-ray occlusion semantics (256-per-ray lists in index order, shared transmittance) for a whole scan grid
        from nlos_gaussian_renderer_b200.occluded import render_transient_batch_occluded
        No, Syo, Sxo, To, Ao = scenes.CONFIGS["cfg2"]
        go = ScanGeometry.wall_grid(Syo, Sxo, A=Ao)
        go = ScanGeometry(go.scan_o.reshape(Syo, Sxo, 3), go.theta_range.reshape(Syo, Sxo, 2), go.phi_range.reshape(Syo, Sxo, 2), Ao, vp_y=go.vp_y, device=dev)
        tao = scenes.time_axis(To)
        po = [p.to(dev).requires_grad_() for p in scenes.make_gaussians(No, 1002)]
        tso = tao["t_samples"].to(dev)
        tf = tb = 0.0
        for it in range(3):
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
            Ho = render_transient_batch_occluded(go, tso, *po, tao["c"], tao["deltaT"])
            e[1].record()
            (Ho * Ho).sum().backward()
            e[2].record()
            torch.cuda.synchronize()
            if it:
                tf += e[0].elapsed_time(e[1]) / 2; tb += e[1].elapsed_time(e[2]) / 2
        line["occlusion"] = {"workload": f"cfg2 with occlusion: {No} Gaussians, {Syo}x{Sxo} scan, {To} bins, A={Ao}, 256-per-ray lists (the reference's cap binds on "
                                         "most rays at this size), numerical integration", "fwd_ms": tf, "bwd_ms": tb,
                             "fwd_includes": "preprocess + per-ray list building + render; bwd: lists + recompute + gradients (nothing of size [rays, T] is stored)",
                             "rays": Syo * Sxo * Ao * Ao, "ray_gaussian_pairs_max": Syo * Sxo * Ao * Ao * 256}
        del Ho, po
        torch.cuda.empty_cache()
    if world == 1 and not args.no_cpu_baseline:
        b = cpu_baseline(args.config, args.cpu_points)
        line["cpu_baseline"] = {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")}
        try:
            line["cpu_dense_torch"] = cpu_dense_torch(args.config)
        except Exception as e:          # (reported baseline only)
            line["cpu_dense_torch"] = {"error": repr(e)}
    print(json.dumps(line))


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
