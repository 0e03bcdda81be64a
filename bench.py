#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native Neuralangelo hot path.

Default workload = BASELINE.json configs[1]: hash-grid encode forward + backward microbenchmark,
16 levels x 8 features, 2^22-entry hashmap (1.46 GB fp32 table), 2^22 points uniform in the unit
ball mapped to [0.25,0.75]^3, centre + 6 finite-difference taps (eps = 1/2048 world = 1/8192 unit),
dOut ~ N(0,1).  One "step" = zero the table gradient, encode 7 x 2^22 rows, scatter 7 x 2^22 rows.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload encode]

Prints ONE JSON line (rank 0).  `value` is whole-job algorithmic GB/s with inputs resident in HBM;
`e2e` is the same metric through the public API with the step's points arriving from pinned HOST
memory and a device->host read of the step's result inside the timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GROWTH = float(np.exp((np.log(2048) - np.log(32)) / 15))  # modules.py:38-41 for min_logres 5, max_logres 11
HG_CFG = (16, 8, 22, 32, GROWTH)
N_POINTS = 1 << 22
EPS_WORLD = 1.0 / 2048
ROW_BYTES = 12 + 128 * 4  # one evaluation: xyz in + one 128-float feature row


def six_taps(eps):
    e = float(np.float32(eps))
    return ((0, 0, 0), (e, 0, 0), (-e, 0, 0), (0, e, 0), (0, -e, 0), (0, 0, e), (0, 0, -e))


def unit_ball_points(n, seed, device):
    """Uniform in the unit ball (radius 1, world coordinates; the [-2,2] volume maps it to [0.25,0.75]^3)."""
    g = torch.Generator(device=device).manual_seed(seed)
    p = torch.randn(n, 3, device=device, generator=g)
    p = p / p.norm(dim=-1, keepdim=True) * torch.rand(n, 1, device=device, generator=g) ** (1.0 / 3.0)
    return p.contiguous()


def algorithmic_bytes(n_points, n_taps, table_bytes):
    fwd = n_points * n_taps * ROW_BYTES + table_bytes
    bwd = n_points * n_taps * ROW_BYTES + 2 * table_bytes
    return fwd, bwd


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return json.load(fh), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class Watchdog:
    """If the guarded block does not finish in `seconds`, print the pending JSON line (rank 0) and leave the process."""

    def __init__(self, seconds, line):
        self.seconds, self.line = seconds, line

    def _fire(self):
        if self.line is not None:
            self.line["extras"] = {"error": f"extras did not finish within {self.seconds} s"}
            print(json.dumps(self.line), flush=True)
        os._exit(0)

    def __enter__(self):
        self.timer = threading.Timer(self.seconds, self._fire)
        self.timer.daemon = True
        self.timer.start()
        return self

    def __exit__(self, *exc):
        self.timer.cancel()
        return False


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            try:
                sm.append(float(r[0])), mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the CPU restatement of the tcnn hash grid on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_encode_sample(n_sample, threads):
    """Runs fwd+bwd of the same workload on `n_sample` points on the CPU (oracle port; tcnn itself is
    CUDA-only and not installable here).  Returns (seconds, equivalent GB/s of the full-size metric)."""
    from oracle import hashgrid_ref as hr  # CPU baseline leg: the one place bench.py executes oracle/
    torch.set_num_threads(threads)
    tbl = hr.level_table(*HG_CFG)
    table = hr.init_table(tbl)
    pts = unit_ball_points(n_sample, 0, "cpu")
    taps = six_taps(EPS_WORLD)
    g = torch.Generator().manual_seed(1)
    dout = torch.randn(len(taps), n_sample, 128, generator=g)
    t0 = time.perf_counter()
    hr.encode_taps(tbl, table, pts, -2.0, 2.0, taps)
    for t, off in enumerate(taps):
        p = pts + torch.tensor(off, dtype=torch.float32)
        hr.encode_backward(tbl, hr.unit_map(p, -2.0, 2.0), dout[t], dtype=torch.float32)
    dt = time.perf_counter() - t0
    fwd, bwd = algorithmic_bytes(N_POINTS, len(taps), table.numel() * 4)
    # the full-size metric scaled by the fraction of points processed (table term scaled alike)
    gbps = (fwd + bwd) * (n_sample / N_POINTS) / dt / 1e9
    return dt, gbps


def run_reference(args, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_sample = 1 << 16
    times, vals = [], []
    for i in range(args.warmup + args.steps):
        dt, gbps = cpu_encode_sample(n_sample, threads)
        if i >= args.warmup:
            times.append(dt), vals.append(gbps)
    value = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": "hashgrid_encode_fwd_bwd_algorithmic_GBps", "value": value, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean(times)) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(),
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": threads, "kind": "port",
                         "sample": f"{n_sample} of {N_POINTS} points x 7 taps fwd+bwd per step, full 1.46 GB table; "
                                   "GB/s = full-size algorithmic bytes x (sample/full) / seconds"},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "host-core throughput of the CPU port; it does not depend on the job size, so at --gpus N (weak scaling: "
                "N x the points) the same figure stands beside the N-GPU aggregate",
    }
    print(json.dumps(line), flush=True)


def workload_config():
    return {"workload": "BASELINE.json configs[1]: hash-grid encode fwd+bwd, 16 levels x 8 features, 2^22-entry "
                        "hashmap, 2^22 points/GPU, centre + 6 taps (eps 1/2048)", "points_per_gpu": N_POINTS,
            "taps": 7, "table_bytes": 45724048 * 8 * 4, "feature_layout": "level-major planes [taps][levels][points][8] fp32",
            "l2_policy": "inputs_exceed_l2 (30 GB of rows + 1.46 GB "
            "table per step vs 126 MB L2)", "parallelism": "points sharded across ranks, no data-path collective"}


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def run_ours(args, rank, local_rank, world):
    from neuralangelo_b200 import _lib, build, hashgrid
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    build.build(verbose=False)
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    desc = hashgrid.make_desc(*HG_CFG)
    taps = six_taps(EPS_WORLD)
    um = _lib.UnitMap(-2.0, 0.25)
    g = torch.Generator(device=dev).manual_seed(1337)
    table = (torch.rand(desc.total_entries * 8, device=dev, generator=g) * 2 - 1) * 1e-4
    pts = unit_ball_points(N_POINTS, rank, dev)
    pts_host = pts.cpu().pin_memory()
    g1 = torch.Generator(device=dev).manual_seed(1 + rank)
    # native feature layout: level-major planes [taps][levels][points][F]
    dout = torch.randn(len(taps), 16, N_POINTS, 8, device=dev, generator=g1)
    out = torch.empty(len(taps), 16, N_POINTS, 8, device=dev)
    dtable = torch.zeros_like(table)
    pts_dev = torch.empty_like(pts)
    result_host = torch.empty(129, dtype=torch.float32).pin_memory()
    fwd_bytes, bwd_bytes = algorithmic_bytes(N_POINTS, len(taps), table.numel() * 4)
    launches = [0]
    bwd_launches = hashgrid.bwd_launch_count(desc, N_POINTS, len(taps), 16)

    def step(x, ev=None):
        dtable.zero_()
        if ev:
            ev[0].record()
        hashgrid.encode_taps_planes(desc, table, x, um, 16, taps, out=out)
        if ev:
            ev[1].record()
        hashgrid.scatter_taps_planes(desc, x, um, 16, taps, dout, dtable)
        if ev:
            ev[2].record()
        launches[0] += 1 + bwd_launches      # one forward launch + the backward's launches (finest levels in two passes)

    # End-to-end form: the points of every step start in pinned HOST memory and the step's result summary ends there.
    # The copies are double-buffered on a copy stream, so step k+1's 50 MB upload runs under step k's kernels; the
    # host blocks once, after the last step's read-back.
    copy_stream = torch.cuda.Stream(device=dev)
    pts_dev2 = [pts_dev, torch.empty_like(pts_dev)]
    result_host2 = [result_host, torch.empty(129, dtype=torch.float32).pin_memory()]
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    def run_e2e(n_steps):
        cur = torch.cuda.current_stream()
        for k in range(n_steps):
            b = k & 1
            with torch.cuda.stream(copy_stream):
                if k >= 2:
                    copy_stream.wait_event(consumed[b])
                pts_dev2[b].copy_(pts_host, non_blocking=True)
                copied[b].record(copy_stream)
            cur.wait_event(copied[b])
            step(pts_dev2[b])
            consumed[b].record(cur)
            res = torch.cat([out[0, :, 0, :].reshape(-1), dtable[:262144].sum().reshape(1)])
            result_host2[b].copy_(res, non_blocking=True)
        cur.synchronize()

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(pts)
    barrier()
    launches[0] = 0
    kernel_events = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        barrier()
        start.record()
        for k in range(args.steps):
            step(pts, kernel_events[k])
        stop.record()
        barrier()
    gpu_launches = launches[0]
    elapsed_ms = start.elapsed_time(stop)
    fwd_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in kernel_events]))
    bwd_ms = float(np.mean([e[1].elapsed_time(e[2]) for e in kernel_events]))

    run_e2e(max(2, args.warmup // 2))
    barrier()
    t0 = time.perf_counter()
    run_e2e(args.steps)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3

    if world > 1:
        t = torch.tensor([elapsed_ms, e2e_ms], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        elapsed_ms, e2e_ms = t.tolist()
    del out, dout, dtable, table, pts_dev, pts_dev2
    torch.cuda.empty_cache()
    if rank != 0:
        if not args.no_extras:
            with Watchdog(args.extras_timeout, None):
                collect_extras(args, rank, world, dev)
        return
    ms_per_step = elapsed_ms / args.steps
    step_bytes = (fwd_bytes + bwd_bytes) * world
    value = step_bytes / (ms_per_step * 1e-3) / 1e9
    e2e_value = step_bytes / (e2e_ms / args.steps * 1e-3) / 1e9
    peaks, peak_kind = measured_peaks()
    dominant = ("hashgrid_axis_bwd_kernel", bwd_bytes, bwd_ms) if bwd_ms >= fwd_ms else \
        ("hashgrid_axis_fwd_kernel", fwd_bytes, fwd_ms)
    achieved = dominant[1] / (dominant[2] * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as fh:
            traffic = json.load(fh).get(dominant[0])
    line = {
        "metric": "hashgrid_encode_fwd_bwd_algorithmic_GBps", "value": value, "unit": "GB/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(),
        "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": N_POINTS * 12,
                "d2h_bytes_per_step": 129 * 4, "ms_per_step": e2e_ms / args.steps,
                "pipelining": "double-buffered pinned->device uploads on a copy stream; one host sync after the last read-back"},
        "gpu_launches": gpu_launches,
        "clocks": clocks.summary(),
        "roofline": {"bound": "hbm", "kernel": dominant[0], "achieved": achieved, "peak": peaks["hbm_gbs"],
                     "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": traffic, "peak_source": peak_kind,
                     "kernel_ms": dominant[2], "algorithmic_bytes": dominant[1],
                     "launches_per_call": {"hashgrid_axis_fwd_kernel": 1, "hashgrid_axis_bwd_kernel": bwd_launches},
                     "note": "kernel_ms / achieved / traffic are per backward (forward) CALL: the backward is "
                             f"{bwd_launches} launches of the same kernel (dense levels interleaved in one, the other coarse "
                             "levels in one, each of the finest hashed levels in two row-range passes), timed and profiled "
                             "as one unit",
                     "other_kernels": {"hashgrid_axis_fwd_kernel": {"ms": fwd_ms, "GBps": fwd_bytes / fwd_ms / 1e6},
                                       "hashgrid_axis_bwd_kernel": {"ms": bwd_ms, "GBps": bwd_bytes / bwd_ms / 1e6}}},
    }
    if not args.no_cpu_baseline and world == 1:
        threads = os.cpu_count() or 1
        n_sample = 1 << 17
        dt, gbps = cpu_encode_sample(n_sample, threads)
        line["cpu_baseline"] = {"value": gbps, "unit": "GB/s", "cores": threads, "kind": "port",
                                "sample": f"{n_sample} of {N_POINTS} points x 7 taps fwd+bwd once ({dt:.1f} s), full "
                                          "1.46 GB table; GB/s = full-size algorithmic bytes x (sample/full) / seconds"}
    # secondary measurements last, under a watchdog: whatever happens there (a hung collective included), the
    # headline line is printed exactly once
    line["extras"] = {}
    if not args.no_extras:
        with Watchdog(args.extras_timeout, line):
            line["extras"] = collect_extras(args, rank, world, dev)
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# secondary measurements carried in "extras" (BASELINE.json configs[2..4]); never allowed to break the main line
# ------------------------------------------------------------------------------------------------
def _look_at(eye):
    eye = torch.tensor(eye, dtype=torch.float32)
    fwd = -eye / eye.norm()
    right = torch.linalg.cross(torch.tensor([0.0, 1.0, 0.0]), fwd)
    right = right / right.norm()
    R = torch.stack([right, torch.linalg.cross(fwd, right), fwd])
    return torch.cat([R, (-R @ eye)[:, None]], dim=1)


def extra_train_step(rank, world, dev, steps=200, warmup=50, taps=4, force_all_levels=True, table_exchange="auto"):
    """BASELINE.json configs[2]/[3]: full training step of the default hash-grid config (16 levels, 2^22, dim 8),
    synthetic cameras (32 poses on a radius-3 sphere, 800x800, focal 1111), random-init weights, 512 rays per
    GPU per step, batch pose/intrinsics/ray indices/target colours arriving from pinned host memory, gradients
    exchanged across ranks (median of >= 200 steps after 50 warm-up).  Returns aggregate rays/s
    and the per-phase device times (forward + backward / gradient exchange + optimizer)."""
    from neuralangelo_b200.config import default_data_config, default_model_config
    from neuralangelo_b200.model import Model
    from neuralangelo_b200.trainer import TrainStep
    torch.manual_seed(0)
    cfg = default_model_config()
    cfg.object.sdf.gradient.taps = taps
    model = Model(cfg, default_data_config()).to(dev)
    ts = TrainStep(model, max_iter=500000, warm_up_end=5000, use_cuda_graph=True, table_exchange=table_exchange)
    if force_all_levels:
        ts.current_iteration = 5000 + 16 * 5000
    g = torch.Generator().manual_seed(1000 + rank)
    poses = []
    for _ in range(32):
        v = torch.randn(3, generator=g)
        poses.append(_look_at((v / v.norm() * 3.0).tolist()))
    intr = torch.tensor([[1111.0, 0, 400.0], [0, 1111.0, 400.0], [0, 0, 1.0]])
    host = [dict(pose=poses[i % 32][None].pin_memory(), intr=intr[None].pin_memory(),
                 idx=torch.tensor([i % 32]).pin_memory(),
                 ray_idx=torch.randint(0, 640000, (1, 512), generator=g).pin_memory(),
                 image_sampled=torch.rand(1, 512, 3, generator=g).pin_memory()) for i in range(8)]

    def step(i):
        data = {k: v.to(dev, non_blocking=True) for k, v in host[i % 8].items()}
        return ts.step(data)

    for i in range(warmup):
        step(i)
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    ts.record_phases, ts.phase_events = True, []
    t0 = time.perf_counter()
    for i in range(steps):
        loss = step(i)
    loss_host = float(loss.detach())   # device->host read of the step's result
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    ev = ts.phase_events
    fb = [e[0].elapsed_time(e[1]) for e in ev]
    ex = [e[1].elapsed_time(e[2]) for e in ev]
    per_step = [a[0].elapsed_time(b[0]) for a, b in zip(ev[:-1], ev[1:])]      # start to start: the steady-state period
    med = statistics.median(per_step) * 1e-3
    if world > 1:
        t = torch.tensor([dt, med, statistics.median(fb), statistics.median(ex)], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        dt, med, fb_ms, ex_ms = t.tolist()
    else:
        fb_ms, ex_ms = statistics.median(fb), statistics.median(ex)
    exchange = "none (1 GPU)" if world == 1 else ("multimem: one kernel (NVLink multicast reduce + AdamW + broadcast)"
                                                  if ts.multimem else "nccl: reduce-scatter / owner AdamW / all-gather")
    del ts, model
    torch.cuda.empty_cache()
    return {"rays_per_s": 512 * world / med, "ms_per_step": med * 1e3, "mean_ms_per_step_host_clock": dt * 1e3,
            "rays_per_gpu": 512, "taps": taps, "active_levels": 16 if force_all_levels else 4, "steps": steps,
            "warmup": warmup, "last_loss": loss_host,
            "phases_ms": {"forward_backward": fb_ms, "exchange_and_optimizer": ex_ms}, "table_exchange": exchange,
            "timing": "median over steps of the device time between consecutive step starts (CUDA events), max over "
                      "ranks; includes H2D of the batch, the gradient exchange and the optimizer"}


def extra_mesh(rank, world, dev, res=2048):
    """BASELINE.json configs[4] at full size: isosurface extraction at resolution 2048, block 128 (4096 blocks, 8.78 G
    lattice points): blocked SDF query of the default hash-grid model (random init) + marching cubes, blocks sharded
    across the ranks of the launch, vertex / face buffers gathered on rank 0 (about 7 s on one B200)."""
    from neuralangelo_b200 import mesh
    from neuralangelo_b200.config import default_model_config
    from neuralangelo_b200.modules import NeuralSDF
    torch.manual_seed(0)
    net = NeuralSDF(default_model_config().object.sdf).to(dev).eval()
    net.warm_up_end = 5000
    net.set_active_levels(500000)
    bounds = np.array([[-1.0, 1.0]] * 3)
    fn = mesh.LatticeSDF(net)      # -sdf (scripts/extract_mesh.py:85), block queries straight from the lattice axes
    mesh.extract_mesh(fn, bounds, intv=2.0 / 128, block_res=128)   # warm-up
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    m = mesh.extract_mesh(fn, bounds, intv=2.0 / res, block_res=128)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        dt = float(t)
    g = mesh.LatticeGrid(bounds, 2.0 / res, 128)
    npts = sum(int(np.prod([len(a) for a in g.block_axes(i)])) for i in range(len(g)))
    out = {"resolution": res, "block_res": 128, "seconds": dt, "lattice_points": npts, "Mpoints_per_s": npts / dt / 1e6,
           "projected_2048_cubed_seconds": dt * (2063.0 ** 3) / npts}
    if m is not None:
        out.update(vertices=len(m.vertices), faces=len(m.faces))
    del net
    torch.cuda.empty_cache()
    return out


def extra_cpu_path():
    """BASELINE.json configs[0] on the host cores (reported baseline, not a target): the CPU restatement of the
    reference's own CPU-runnable path -- Fourier-encoded NeuralSDF, 1024 rays x (64 + 32) samples forward + backward,
    then extract_mesh at resolution 256 / block 128 with the serial C marching cubes (oracle/cpu_path.py, pinned to
    the reference by tests/test_oracle_cpu_path.py)."""
    from oracle import cpu_path   # CPU baseline leg
    threads = os.cpu_count() or 1
    out = cpu_path.timed_baseline(n_rays=1024, resolution=256, threads=threads)
    out["kind"] = "port"
    out["cores"] = threads
    return out


def extra_sdf_mlp(dev, m=1 << 21):
    """The only dense contraction of the path: SDF-only evaluation (131 -> 256 -> SDF head) of 2^21 rows on the
    pipelined tcgen05 kernel, input as chunk planes (the hash-grid kernels' layout) and as [M,131] rows."""
    from neuralangelo_b200.mlp import MLPforNeuralSDF, get_activation
    from neuralangelo_b200.fused_mlp import PackedSdfMlp
    torch.manual_seed(0)
    mlp = MLPforNeuralSDF([131, 256, 256], activ=get_activation("softplus", beta=100.0), use_weightnorm=True,
                          geometric_init=True, out_bias=0.5).to(dev)
    packed = PackedSdfMlp(mlp, 100.0)
    planes = torch.randn(17, m, 8, device=dev)
    rows = torch.randn(m, 131, device=dev)

    def timed(fn, iters=5):
        for _ in range(2):
            fn()
        ts = []
        for _ in range(iters):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    ms_p = timed(lambda: packed.forward_planes(planes, m))
    ms_r = timed(lambda: packed.forward(rows, with_sdf=True, with_feat=False))
    mma_flop = 3 * 2 * m * 144 * 256    # three fp16 MMAs per k-step (split-precision operands), K padded to 144
    peaks = measured_peaks()[0]
    return {"mma_frac_of_measured_dense_bf16_peak": mma_flop / ms_p / 1e9 / peaks["bf16_tflops"],
            "rows": m, "planes_ms": ms_p, "planes_Grows_per_s": m / ms_p / 1e6, "rows_layout_ms": ms_r,
            "mma_TFLOPs_planes": mma_flop / ms_p / 1e9, "fp32_equiv_TFLOPs_planes": 2 * m * (131 * 256 + 256) / ms_p / 1e9}


def extra_composite(dev, log2_rays=20):
    """Compositing microbenchmark: the fused NeuS renderer (alphas -> transmittance -> weights ->
    composites, object + background samples) forward and backward on 2^20 rays x (128 + 32) samples; achieved GB/s of
    ALGORITHMIC bytes and the fraction of the measured HBM peak."""
    from neuralangelo_b200 import render
    R, N, NB = 1 << log2_rays, 128, 32
    g = torch.Generator(device=dev).manual_seed(0)
    ray = torch.nn.functional.normalize(torch.randn(1, R, 3, device=dev, generator=g), dim=-1)
    dists = torch.sort(torch.rand(1, R, N, 1, device=dev, generator=g) * 2 + 1, dim=2).values
    sdf = ((2.0 - dists) * 0.4 + 0.02 * torch.randn(1, R, N, 1, device=dev, generator=g)).requires_grad_(True)
    grad = (-ray[:, :, None, :] + 0.3 * torch.randn(1, R, N, 3, device=dev, generator=g)).requires_grad_(True)
    rgb = torch.rand(1, R, N, 3, device=dev, generator=g).requires_grad_(True)
    dens = torch.rand(1, R, NB, device=dev, generator=g).requires_grad_(True)
    dbg = torch.sort(torch.rand(1, R, NB, 1, device=dev, generator=g) * 10 + 3.2, dim=2).values
    rgbb = torch.rand(1, R, NB, 3, device=dev, generator=g).requires_grad_(True)
    far = dists[:, :, -1:, :] + 0.1
    s_var = torch.tensor(3.0, device=dev, requires_grad=True)

    def timed(fn, iters=5):
        for _ in range(2):
            fn()
        ts = []
        for _ in range(iters):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    kw = dict(dist_far=far, densities_bg=dens, dists_bg=dbg, rgbs_bg=rgbb, want_weights=False)

    def fwd():
        with torch.no_grad():
            render.neus_render(ray, sdf, grad, dists, s_var, rgb, **kw)

    peak = measured_peaks()[0]["hbm_gbs"]
    b_in = N * 32 + NB * 20 + 12 + 4          # per ray: sdf 4 + grad 12 + dist 4 + rgb 12 per object sample; 20 per bg sample
    ms_f = timed(fwd)
    by_f = R * (b_in + 16)
    out, _, _ = render.neus_render(ray, sdf, grad, dists, s_var, rgb, **kw)
    d_rgb = torch.randn_like(out)
    ms_b = timed(lambda: torch.autograd.grad(out, [sdf, grad, rgb, dens, rgbb, s_var], d_rgb, retain_graph=True))
    by_b = R * (b_in + 12 + N * 28 + NB * 16)  # inputs re-read + d_rgb; d_sdf, d_grad, d_rgb, d_dens, d_rgb_bg written
    kw_w = dict(kw, want_weights=True)

    def fwd_w():
        with torch.no_grad():
            render.neus_render(ray, sdf, grad, dists, s_var, rgb, **kw_w)
    ms_fw = timed(fwd_w)
    by_fw = by_f + R * (N + NB) * 4     # + the compositing weights the training step keeps (model.py:141)
    return {"rays": R, "samples": [N, NB],
            "fwd": {"ms": ms_f, "GBps": by_f / ms_f / 1e6, "frac_of_hbm_peak": by_f / ms_f / 1e6 / peak},
            "fwd_with_weights": {"ms": ms_fw, "GBps": by_fw / ms_fw / 1e6, "frac_of_hbm_peak": by_fw / ms_fw / 1e6 / peak},
            "bwd": {"ms": ms_b, "GBps": by_b / ms_b / 1e6, "frac_of_hbm_peak": by_b / ms_b / 1e6 / peak}}


def collect_extras(args, rank, world, dev):
    extras = {}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            extras["cpu_path_configs0"] = extra_cpu_path()
        except Exception as e:  # noqa: BLE001
            extras["cpu_path_configs0"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    for name, fn in (("sdf_mlp", lambda: extra_sdf_mlp(dev)),
                     ("composite", lambda: extra_composite(dev)),
                     ("train_step", lambda: extra_train_step(rank, world, dev)),
                     ("train_step_taps6", lambda: extra_train_step(rank, world, dev, taps=6)),
                     ("train_step_4_active_levels", lambda: extra_train_step(rank, world, dev, force_all_levels=False)),
                     ("mesh_extract", lambda: extra_mesh(rank, world, dev))):
        try:
            extras[name] = fn()
        except Exception as e:  # noqa: BLE001 -- extras must never break the headline line
            extras[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
    if world > 1:   # the library-collective form of the same step, for comparison with the fused multicast kernel
        try:
            extras["train_step_nccl_exchange"] = extra_train_step(rank, world, dev, table_exchange="nccl")
        except Exception as e:  # noqa: BLE001
            extras["train_step_nccl_exchange"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    return extras


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="encode", choices=["encode"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary train-step / mesh measurements")
    ap.add_argument("--extras-timeout", type=float, default=240.0)
    ap.add_argument("--run-timeout", type=float, default=540.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    # whole-run guard: a rank that never shows up (or a wedged collective) must not hang the launch for ever
    def _give_up():
        if rank == 0:
            print(json.dumps({"metric": "hashgrid_encode_fwd_bwd_algorithmic_GBps", "value": None, "unit": "GB/s",
                              "n_gpus": world, "error": f"bench.py did not finish within {args.run_timeout} s"}), flush=True)
        os._exit(3)
    guard = threading.Timer(args.run_timeout, _give_up)
    guard.daemon = True
    guard.start()
    if world > 1:
        import datetime
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local_rank),
                                             timeout=datetime.timedelta(seconds=180))
    try:
        run_ours(args, rank, local_rank, world)
    finally:
        guard.cancel()
        if world > 1:
            torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
