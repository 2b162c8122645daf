#!/usr/bin/env python
"""bench.py -- headline benchmark of the geometric segmentation hot path.

Metric (BASELINE.json): Mpoints/s segmented (normals + region-grow + plane fit).
Workload at N=1: config C2 -- a batch of 256 synthetic 640x480 organized depth
frames (floor + 3 boxes + noise), plane + cluster segmentation, frames sharded
frame-wise (no collective; weak scaling: 256 frames per GPU).

One step = one pass of the pipeline (ComputeNormalsOrganized ->
SegmentRegions<PlanarRegion> -> SegmentRegions<ClusterRegion>) over the batch.
  value : inputs resident in HBM, device time (CUDA events), max over ranks
  e2e   : same call with HOST (pinned) buffers: H2D of the points and D2H of
          normals + labels inside the timed region
  --impl reference : the reference algorithm's CPU implementation (oracle/_ref
          when built, else the oracle port) on all host cores, bounded sample.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

# Four workspaces x three streams run concurrently in the e2e arm; with the default 8 hardware
# queues independent streams share a queue and serialise (measured: 46 -> 35 ms per step).
# Must be set before the CUDA context exists.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200")
for _p in (PKG, os.path.join(ROOT, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np

ROWS, COLS = 480, 640
N_PTS = ROWS * COLS
FRAMES_PER_GPU = 256
METRIC = "Mpoints/s segmented (normals+region-grow+plane fit)"
UNIT = "Mpoints/s"


def _gen_frame(i):
    """Frame id i = replica * 256 + k: frame k of the C2 sweep with the replica's noise seeds."""
    import synth
    return synth.c2_frame(i % FRAMES_PER_GPU, FRAMES_PER_GPU, ROWS, COLS, replica=i // FRAMES_PER_GPU)


def gen_frames(indices):
    """C2 frames, generated in parallel on the host cores."""
    import multiprocessing as mp
    procs = max(1, min(len(indices), len(os.sched_getaffinity(0)) or 1, 32))
    if procs == 1:
        return np.stack([_gen_frame(i) for i in indices])
    with mp.get_context("fork").Pool(procs) as pool:
        return np.stack(pool.map(_gen_frame, list(indices), chunksize=1))


def cpu_kind():
    """"reference": oracle/_ref/libref.so (the reference's own .h/.cc compiled in place against
    shim headers, built where /root/reference exists and shipped prebuilt); else the "port"."""
    import oracle
    return "reference" if oracle.ref_available() else "port"


def _cpu_pipeline(xyz, want_regions=False):
    """The reference's CPU path for one frame: ComputeNormalsOrganized ->
    SegmentRegions<PlanarRegion> -> SegmentRegions<ClusterRegion> (oracle/_ref when built,
    else the oracle port)."""
    import contextlib
    import oracle
    with (oracle.reference() if oracle.ref_available() else contextlib.nullcontext()):
        nrm = oracle.normals_organized(xyz, ROWS, COLS)
        seeds = np.arange(N_PTS - 1, -1, -1, dtype=np.int32)
        labels = np.full(N_PTS, -1, np.int32)
        lab, regs = oracle.segment_planar(xyz, nrm, ROWS, COLS, seeds, labels)
        lab2, cl = oracle.segment_cluster(xyz, ROWS, COLS, seeds, lab, id_offset=regs.n)
    if want_regions:
        return lab2, regs, cl
    return lab2, regs.n, cl.n


_FRAMES = None  # shared with forked workers (copy-on-write)


def _cpu_worker(i):
    _, n_planar, n_cluster = _cpu_pipeline(_FRAMES[i])
    return n_planar, n_cluster


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for k, nme in enumerate(names):
                    if r[2 + k].lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def _parse_cpulist(txt):
    cpus = set()
    for part in txt.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def bind_to_gpu_numa_node(local_rank):
    """Pins this rank to the CPUs of its GPU's NUMA node (when the container's CPU set has any),
    so that the pinned host buffers are first-touched on the node the GPU's PCIe root hangs off
    and the H2D / D2H copies of the e2e arm do not cross the socket interconnect.  Returns a
    short description for the JSON line."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as f:
            node = int(f.read())
        if node < 0:
            return "gpu %s: no NUMA node reported" % bdf
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            want = _parse_cpulist(f.read()) & os.sched_getaffinity(0)
        if not want:
            return "gpu %s on node %d: none of its CPUs in this container's CPU set" % (bdf, node)
        os.sched_setaffinity(0, want)
        return "gpu %s on node %d: bound to %d CPUs" % (bdf, node, len(want))
    except Exception as e:  # no sysfs / attribute: run unbound
        return "unbound (%s)" % type(e).__name__


def ncu_traffic():
    """DRAM bytes per launch / warp instructions per point of the kernels from the committed ncu
    capture of this same command (profiles/ncu_traffic.json, written from `ncu --set full`)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def run_reference(args, rank, world):
    """CPU arm: reference algorithm on all host cores, bounded sample per step."""
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    sample = min(FRAMES_PER_GPU, int(os.environ.get("PCSEG_BENCH_REF_SAMPLE", 4 * cores)))  # frames per step
    kind = cpu_kind()
    if kind == "reference":
        import oracle
        with oracle.reference():
            oracle.lib()   # dlopen oracle/_ref/libref.so in THIS process: the forked workers inherit the mapping
    global _FRAMES
    _FRAMES = gen_frames(list(range(sample)))
    with mp.get_context("fork").Pool(cores) as pool:
        times = []
        for it in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            pool.map(_cpu_worker, list(range(sample)), chunksize=1)
            dt = time.perf_counter() - t0
            if it >= args.warmup:
                times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    value = sample * N_PTS / (ms * 1e-3) / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic (seeded analytic ray casting, SURVEY.md 8d C2)",
        "config": {"workload": "C2: 640x480 organized depth frames, normals+planes+clusters, seeds all-descending",
                   "frames_per_step": sample, "rows": ROWS, "cols": COLS},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": "%d frames per step over %d worker processes" % (sample, cores)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_c5(args, rank, local_rank, world):
    """--config c5: SlidingMeanShift of ONE 10 M-point cloud.  The only path with a
    real exchange step: cloud + labels replicated, the seed cells of the iteration
    phase sharded by contiguous linear-index ranges over the ranks, mode records all-gathered with
    NCCL, sort + growth once on rank 0.  Total work is fixed: strong scaling."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    import parallel
    import pcseg
    import synth
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    rows, cols, iters = args.c5_rows, args.c5_cols, 5
    n = rows * cols
    xyz = synth.c5_full(rows, cols)                   # the same seeded cloud on every rank
    host_xyz = torch.from_numpy(xyz).pin_memory()
    host_lab = torch.full((n,), -1, dtype=torch.int32).pin_memory()
    d_xyz = host_xyz.cuda()
    d_lab = host_lab.cuda()
    out_lab = torch.empty((n,), dtype=torch.int32).pin_memory()
    ctx = pcseg.Context(local_rank)
    n_seeds = int((~np.isnan(xyz[:, 0])).sum())
    state = {}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def step(e2e):
        if e2e:   # the cloud and the labels come from (pinned) host memory every step
            d_xyz.copy_(host_xyz, non_blocking=True)
            d_lab.copy_(host_lab, non_blocking=True)
        t0 = time.perf_counter()
        labels, regs = parallel.sharded_mean_shift(ctx, d_xyz, rows, cols, d_lab, rank, world, iterations=iters,
                                                   labels_out=out_lab if e2e else None, labels_on_device=not e2e)
        state["t_ms"] = (time.perf_counter() - t0) * 1e3
        if rank == 0:
            state["regions"], state["labels"] = regs, labels

    def timed(e2e, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            step(e2e)
        e1.record()
        e1.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        barrier()
        # the library calls block the host until their stream is idle, so the events bracket them
        return max_over_ranks(max(e0.elapsed_time(e1), wall)) / steps

    for _ in range(max(args.warmup, 3)):
        step(False)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    ms = timed(False, args.steps)
    launches = ctx.launches - l0
    clocks = sampler.stop() if rank == 0 else None
    step(True)
    e2e_ms = timed(True, args.steps)
    # iteration phase alone (modes kernel + all-gather)
    def modes_fn(b, e, out):
        ctx._check(ctx.L.pcseg_mean_shift_modes_range(ctx.h, C.c_void_p(d_xyz.data_ptr()), rows, cols,
                                                      C.c_void_p(d_lab.data_ptr()), iters, b, e, C.c_void_p(out.data_ptr())))
    it_ms = []
    for _ in range(5):
        barrier()
        t0 = time.perf_counter()
        parallel.sharded_mean_shift_modes(modes_fn, n, rank, world, device=d_xyz.device)
        torch.cuda.synchronize()
        it_ms.append(max_over_ranks((time.perf_counter() - t0) * 1e3))
    it_ms = statistics.median(it_ms)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    regs = state["regions"]
    # CPU baseline + parity on a bounded sample: a crop of whole tiles of the same cloud
    cr, cc = min(rows, 5 * max(rows // 20, 16)), min(cols, 5 * max(cols // 20, 16))
    crop = np.ascontiguousarray(xyz.reshape(cols, rows, 3)[:cc, :cr].reshape(-1, 3))
    import oracle
    t0 = time.perf_counter()
    o_lab, o_r, _ = oracle.mean_shift(crop, cr, cc, np.full(cr * cc, -1, np.int32), iterations=iters)
    cpu_s = time.perf_counter() - t0
    g_lab, g_r = ctx.mean_shift(crop, cr, cc, np.full(cr * cc, -1, np.int32), iterations=iters)
    same = bool(np.array_equal(o_lab, g_lab)) and o_r.n == g_r.n and \
        all(np.array_equal(o_r.inliers[k], g_r.inliers[k]) for k in range(min(o_r.n, g_r.n)))
    peak, peak_kind = measured_peak_gbs()
    alg = 44.0 * n_seeds   # 12 B point + 4 B label read, 28 B mode record written per seed
    line = {
        "metric": METRIC, "value": n / (ms * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic (SURVEY.md 8d C5: blobs on a tiled ground sheet, seeded)",
        "config": {"workload": "C5: SlidingMeanShift of one %dx%d = %.1f M point cloud, %d iterations; seed cells sharded over "
                               "the ranks, mode records all-gathered (NCCL), sort + growth on rank 0" % (rows, cols, n / 1e6, iters),
                   "rows": rows, "cols": cols, "seeds": n_seeds, "regions": int(regs.n),
                   "l2": "inputs larger than L2 (%.0f MB of points)" % (n * 12 / 1e6)},
        "iteration_phase_ms": it_ms, "allgather_bytes_per_rank": int(-(-n // world) * 32) if world > 1 else 0,
        "sort_and_growth_ms": ms - it_ms,
        "e2e": {"value": n / (e2e_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": int(n * 16) * world, "d2h_bytes_per_step": int(n * 4),
                "outputs": "label grid to pinned host memory on rank 0 (+ region records)"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "meanshift_modes_kernel (iteration phase incl. the all-gather, max over ranks)",
                     "achieved": alg / (it_ms * 1e-3) / 1e9, "peak": peak * world, "unit": "GB/s",
                     "frac": alg / (it_ms * 1e-3) / 1e9 / (peak * world), "traffic": None,
                     "algorithmic_bytes_per_launch": alg, "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs) x n_gpus",
                     "note": "bound by L1 gathers (121 window cells x 16 B x 5 iterations per seed), SURVEY.md 8d"},
        "cpu_baseline": {"value": cr * cc / cpu_s / 1e6, "unit": UNIT, "cores": 1, "kind": "port",
                         "sample": "a %dx%d crop (25 whole tiles) of the same cloud, single-threaded, %.1f s" % (cr, cc, cpu_s)},
        "parity": {"crop_labels_and_inlier_lists_equal_oracle": same, "crop_regions": int(o_r.n)},
        "clocks": clocks, "host_cores": os.cpu_count(),
    }
    print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=FRAMES_PER_GPU)
    ap.add_argument("--cpu-sample", type=int, default=16)
    ap.add_argument("--seed-mode", default="alldesc", choices=["alldesc", "planesupport"])
    ap.add_argument("--depth", type=int, default=4, choices=[1, 2, 3, 4, 6, 8],
                    help="batch workspaces in flight per GPU (2: pcseg_batch_submit/wait, the copies and the "
                         "normals stage of one batch overlap the growth stages of the other; 1: synchronous)")
    ap.add_argument("--config", default="c2", choices=["c2", "c5"],
                    help="c2 (default): the headline frame-batch workload; c5: mean shift of one 10 M-point cloud, "
                         "seed cells sharded over the ranks + NCCL all-gather (the only collective on the path)")
    ap.add_argument("--c5-rows", type=int, default=2500)
    ap.add_argument("--c5-cols", type=int, default=4000)
    ap.add_argument("--secondary", action="store_true",
                    help="instead of the bench line: one-frame timings of configs C3, C4, C5 with the CPU "
                         "checker timed beside them (tools/bench_configs.py)")
    args = ap.parse_args()
    if args.secondary:
        import runpy
        import oracle
        runpy.run_path(os.path.join(ROOT, "tools", "bench_configs.py"), init_globals={"CPU_BASELINE_MODULE": oracle},
                       run_name="__main__")
        return
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.config == "c5":
        run_c5(args, rank, local_rank, world)
        return

    import torch
    import torch.distributed as dist
    import pcseg

    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_numa_node(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    F = args.frames
    # frame-wise sharding, weak scaling: every rank segments its own 256-frame C2 batch (the same
    # camera sweep, rank-specific noise seeds), no data-path collective
    ids = [rank * FRAMES_PER_GPU + k for k in range(F)]
    host = torch.from_numpy(gen_frames(ids)).pin_memory()           # [F, N, 3] pinned
    dev_xyz = host.to("cuda", non_blocking=False)

    pinned_keep = []

    def pinned(shape, dtype):
        """numpy view of pinned host memory (kept alive in pinned_keep)."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        t = torch.empty(max(n, 1), dtype=torch.uint8).pin_memory()
        pinned_keep.append(t)
        return t.numpy()[:n].view(dtype).reshape(shape)

    ctx = pcseg.Context(local_rank)
    D = args.depth
    batches = [pcseg.Batch(ctx, F, ROWS, COLS) for _ in range(D)]
    batch = batches[0]
    # one pinned output set per workspace: label grid + region results (records, inlier / boundary /
    # discontinuity lists, hulls); a host that cannot pin that much runs with fewer in flight
    ol_np, res, res_do, on_np = [], [], [], None
    try:
        for _ in range(D):
            ol_np.append(pinned((F, N_PTS), np.int32))
            res.append(pcseg.BatchResults(F, N_PTS, hull=True, alloc=pinned))
    except RuntimeError:
        D = min(len(ol_np), len(res))
        if D == 0:
            raise
        ol_np, res = ol_np[:D], res[:D]
        for b in batches[D:]:
            b.close()
        batches = batches[:D]
    mode = pcseg.SEEDS_ALL_DESCENDING if args.seed_mode == "alldesc" else pcseg.SEEDS_PLANE_SUPPORT
    params = pcseg.pipeline_params(seed_mode=mode, run_cluster=True)
    params_ps = pcseg.pipeline_params(seed_mode=pcseg.SEEDS_PLANE_SUPPORT, run_cluster=True)
    stream = torch.cuda.ExternalStream(batch.stream)
    # device-resident arm: same job (growth + region pack with hulls), results stay in HBM
    res_dev = [pcseg.BatchResults(F, N_PTS, hull=True, alloc=lambda shape, dtype: None) for _ in range(D)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    stage_acc = {"normals": 0.0, "seeds": 0.0, "planar": 0.0, "cluster": 0.0, "pack": 0.0}

    def run_steps(steps, submit, collect=False, nws=None):
        """`steps` passes of the pipeline, step i on workspace i % D; a workspace is waited for
        (and its stage timers read) before it is reused, and all are drained at the end."""
        nws = nws or D
        busy = [False] * D
        for i in range(steps):
            j = i % nws
            if busy[j]:
                batches[j].wait()
                if collect:
                    for k, v in batches[j].stage_times().items():
                        stage_acc[k] += v
            submit(j)
            busy[j] = True
        for j in range(D):
            if busy[j]:
                batches[j].wait()
                if collect:
                    for k, v in batches[j].stage_times().items():
                        stage_acc[k] += v

    def timed(submit, steps, collect=False, nws=None):
        """K steps bracketed by barrier+sync; device time between two CUDA events on the first
        workspace's stream (idle at both records, so they bracket the whole region); also the
        host wall clock."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        run_steps(steps, submit, collect, nws)
        e1.record(stream)
        e1.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)), max_over_ranks(wall)

    def timed3(submit, steps, host_clock, collect=False, nws=None):
        """Three timed regions of K steps: the median one is reported, min/max as the spread (the
        D-workspace schedule is dynamic and moves by a few per cent run to run)."""
        ms = []
        for rep in range(3):
            d, w = timed(submit, steps, collect and rep == 0, nws)
            ms.append((max(d, w) if host_clock else d) / steps)
        ms.sort()
        return ms[1], ms

    total_pts = F * N_PTS * world

    def mpts(ms_step):
        return total_pts / (ms_step * 1e-3) / 1e6

    # ---- device-resident arm (value) ----
    submit_dev = lambda j: batches[j].submit_regions(dev_xyz.data_ptr(), params, res_dev[j])
    run_steps(max(args.warmup, D), submit_dev)   # every workspace allocates its buffers on first use
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    ms_per_step, dev_spread = timed3(submit_dev, args.steps, host_clock=False, collect=True)
    launches = (ctx.launches - l0) // 3
    clocks = sampler.stop() if rank == 0 else None
    stats = batch.frame_stats()
    value = mpts(ms_per_step)

    # ---- end-to-end arm (host buffers through the C ABI): every step copies its 256 frames in from
    # pinned host memory and copies OUT what the reference's SegmentRegions calls return: the label
    # grid and every frame's PlanarRegion / ClusterRegion (records with plane, centroid, area,
    # estimator sums; inlier, boundary and discontinuity index lists; convex hulls). ----
    host_np = host.numpy()
    submit_e2e = lambda j: batches[j].submit_regions(host_np, params, res[j], labels_out=ol_np[j])
    run_steps(max(2, D), submit_e2e)             # staging buffers of every workspace exist before timing
    e2e_ms_step, e2e_spread = timed3(submit_e2e, args.steps, host_clock=True)
    e2e_value = mpts(e2e_ms_step)
    e2e_d2h = int(F * N_PTS * 4 + res[0].d2h_bytes) * world
    e2e_ok = int(stats[:, 5].max()) == 0 and res[0].s.error_flags == 0
    regions_total = (int(res[0].s.n_planar), int(res[0].s.n_cluster))
    variants = {}
    # variant: label grid only (round 1's e2e)
    submit_lab = lambda j: batches[j].submit(host_np, params, labels_out=ol_np[j])
    run_steps(max(2, D), submit_lab)
    v_ms, v_sp = timed3(submit_lab, args.steps, host_clock=True)
    variants["labels_only"] = {"value": mpts(v_ms), "ms_per_step": v_ms, "d2h_bytes_per_step": int(F * N_PTS * 4) * world,
                               "outputs": "label grid only"}
    # variant: the primary seed list of SURVEY.md 8d, FindSeedPointsFromPlaneSupport on the device
    if args.seed_mode == "alldesc":
        submit_ps = lambda j: batches[j].submit_regions(host_np, params_ps, res[j], labels_out=ol_np[j])
        run_steps(max(2, D), submit_ps)
        v_ms, v_sp = timed3(submit_ps, args.steps, host_clock=True)
        st_ps = batch.stage_times()
        variants["seeds_planesupport"] = {"value": mpts(v_ms), "ms_per_step": v_ms, "spread_ms": v_sp,
                                          "regions": [int(res[0].s.n_planar), int(res[0].s.n_cluster)],
                                          "outputs": "labels + regions, planar seeds = FindSeedPointsFromPlaneSupport (device)",
                                          "stage_ms_last_workspace_overlapped": st_ps}
        submit_dps = lambda j: batches[j].submit_regions(dev_xyz.data_ptr(), params_ps, res_dev[j])
        run_steps(D, submit_dps)
        v_ms, v_sp = timed3(submit_dps, args.steps, host_clock=False)
        variants["seeds_planesupport"]["value_device_resident"] = mpts(v_ms)
        batch.run(dev_xyz.data_ptr(), params_ps)
        variants["seeds_planesupport"]["stage_ms"] = batch.stage_times()
    # variant: DetectedObjects payload as well (inlier + boundary xyz gathers of ps/detected_objects.cc:21-48);
    # two workspaces in flight (the xyz result buffers are 1.9 GB of pinned memory each)
    D2 = min(D, 2)
    try:
        keep0 = len(pinned_keep)
        res_do = [pcseg.BatchResults(F, N_PTS, inlier_xyz=True, boundary_xyz=True, hull=True, alloc=pinned, share=res[j])
                  for j in range(D2)]
        submit_do = lambda j: batches[j].submit_regions(host_np, params, res_do[j], labels_out=ol_np[j])
        run_steps(D2, submit_do, nws=D2)
        v_ms, v_sp = timed3(submit_do, args.steps, host_clock=True, nws=D2)
        variants["detected_objects"] = {"value": mpts(v_ms), "ms_per_step": v_ms, "spread_ms": v_sp,
                                        "d2h_bytes_per_step": int(F * N_PTS * 4 + res_do[0].d2h_bytes) * world,
                                        "workspaces_in_flight": D2,
                                        "outputs": "labels + regions + inlier/boundary xyz (DetectedObjects geometry)"}
    except RuntimeError as e:
        variants["detected_objects"] = {"skipped": "host could not pin the xyz result buffers (%s)" % type(e).__name__}
    res_do = None
    del pinned_keep[keep0:]
    # variant: normals copied back too
    try:
        on_np = [pinned((F, N_PTS, 3), np.float32) for _ in range(D2)]
        submit_n = lambda j: batches[j].submit_regions(host_np, params, res[j], normals_out=on_np[j], labels_out=ol_np[j])
        run_steps(D2, submit_n, nws=D2)
        v_ms, v_sp = timed3(submit_n, args.steps, host_clock=True, nws=D2)
        variants["with_normals"] = {"value": mpts(v_ms), "ms_per_step": v_ms, "workspaces_in_flight": D2,
                                    "d2h_bytes_per_step": e2e_d2h + int(F * N_PTS * 12) * world,
                                    "outputs": "labels + regions + normals"}
    except RuntimeError as e:
        variants["with_normals"] = {"skipped": "host could not pin the normals buffers (%s)" % type(e).__name__}
    on_np = None
    del pinned_keep[keep0:]

    # ---- host <-> device path of this box: plain pinned copies, both directions at once, all ranks
    # together (the ceiling of any e2e number; measured, not assumed) ----
    probe_bytes = 256 << 20
    ph = torch.empty(probe_bytes, dtype=torch.uint8).pin_memory()
    ph2 = torch.empty(probe_bytes, dtype=torch.uint8).pin_memory()
    pd = torch.empty(probe_bytes, dtype=torch.uint8, device="cuda")
    pd2 = torch.empty(probe_bytes, dtype=torch.uint8, device="cuda")
    s_a, s_b = torch.cuda.Stream(), torch.cuda.Stream()
    barrier()
    t0 = time.perf_counter()
    for _ in range(4):
        with torch.cuda.stream(s_a):
            pd.copy_(ph, non_blocking=True)
        with torch.cuda.stream(s_b):
            ph2.copy_(pd2, non_blocking=True)
    torch.cuda.synchronize()
    probe_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    host_copy_gbs = world * 4 * 2 * probe_bytes / probe_s / 1e9
    e2e_copy_gbs = (int(F * N_PTS * 12) * world + e2e_d2h) / (e2e_ms_step * 1e-3) / 1e9
    del ph, ph2, pd, pd2

    ol0 = ol_np[0]
    run_steps(1, submit_e2e)   # workspace 0 holds the headline configuration's result again (parity check below)

    # ---- one batch alone, synchronous: its latency (all 256 frames submitted up front; a frame's
    # result is complete when its batch is, so this is every frame's completion - submit time) ----
    sync_ms = []
    for it in range(3):
        t0 = time.perf_counter()
        batch.submit_regions(dev_xyz.data_ptr(), params, res_dev[0])
        batch.wait()
        sync_ms.append((time.perf_counter() - t0) * 1e3)
    batch_latency_ms = statistics.median(sync_ms)
    stage_ms_sync = batch.stage_times()   # one batch alone: no other workspace's kernels beside it

    # ---- single frame latency ----
    b1 = pcseg.Batch(ctx, 1, ROWS, COLS)
    lat = []
    for it in range(5):
        t0 = time.perf_counter()
        b1.run(dev_xyz.data_ptr(), params)
        lat.append((time.perf_counter() - t0) * 1e3)
    b1.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel.  Kernel durations are the CUDA-event stage timings of
    # one batch run ALONE (synchronous pass above): in the timed region D workspaces overlap, so a
    # stage's elapsed time there (stage_ms) also contains the other workspaces' kernels. ----
    stage_ms_overlapped = {k: v / args.steps for k, v in stage_acc.items()}
    stage_ms = stage_ms_sync
    peak, peak_kind = measured_peak_gbs()
    dom = max(("planar", "normals", "cluster"), key=lambda k: stage_ms[k])
    fused = stage_ms["cluster"] < 0.05   # cluster stage runs inside planar_grow_kernel (per-frame fusion)
    alg_bytes = {"normals": 24.0 * F * N_PTS,            # 12 B read + 12 B write per point
                 "planar": 24.0 * F * N_PTS,             # point + label rd/wr + inlier index per point entering
                 "cluster": 24.0 * float((F * N_PTS) - int(stats[:, 2].sum()))}
    if fused:
        alg_bytes["planar"] += alg_bytes["cluster"]
    traffic = ncu_traffic()
    roof = {}
    for k in ("normals", "planar") if fused else ("normals", "planar", "cluster"):
        ach = alg_bytes[k] / (stage_ms[k] * 1e-3) / 1e9 if stage_ms[k] > 0 else 0.0
        kname = {"planar": "planar_grow_kernel", "normals": "normals_kernel", "cluster": "cluster_grow_kernel"}[k]
        tr = traffic.get(kname, {})
        roof[k] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                   "traffic": tr.get("dram_bytes_per_launch") if F == FRAMES_PER_GPU else None,
                   "algorithmic_bytes_per_launch": alg_bytes[k],
                   "kernel_ms": stage_ms[k], "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)"}
        if "warp_instructions_per_point" in tr:
            # second roof for an issue-bound kernel: warp instructions issued against 148 SMs x 4
            # schedulers x 1 instruction / cycle at the sampled SM clock
            mhz = (clocks or {}).get("sm_mhz") or 1965.0
            issue_peak = 148 * 4 * mhz * 1e6
            wi = tr["warp_instructions_per_point"] * F * N_PTS / 32.0
            roof[k]["issue"] = {"bound": "issue", "warp_instructions_per_launch": wi,
                                "achieved": wi / (stage_ms[k] * 1e-3) / 1e9, "peak": issue_peak / 1e9,
                                "unit": "G warp-instr/s", "frac": wi / (stage_ms[k] * 1e-3) / issue_peak,
                                "source": tr.get("source")}
    roofline = dict(roof[dom])
    roofline["kernel"] = {"planar": "planar_grow_kernel" + (" (planar + fused cluster growth)" if fused else ""),
                          "normals": "normals_kernel", "cluster": "cluster_grow_kernel"}[dom]
    roofline["traffic_source"] = traffic.get("source")
    roofline["note"] = ("kernel_ms: one batch alone (CUDA events on its stream); latency-bound serial FIFO chain "
                        "(one re-fit per 30 inliers), not bandwidth-bound; see DESIGN.md section 8 and profiles/")

    # ---- CPU baseline: the reference's own sources (oracle/_ref), single thread, bounded sample;
    # the same run is the parity check of the e2e outputs (labels + region inlier lists) ----
    ns = min(args.cpu_sample, F)
    t0 = time.perf_counter()
    mismatched_frames = 0
    for i in range(ns):
        lab_cpu, pr, cr = _cpu_pipeline(host_np[i], want_regions=True)
        g = res[0].frame_regions(i, "planar")
        c = res[0].frame_regions(i, "cluster")
        same = np.array_equal(lab_cpu, ol0[i]) and g.n == pr.n and c.n == cr.n
        same = same and all(np.array_equal(g.inliers[k], pr.inliers[k]) for k in range(min(g.n, pr.n)))
        same = same and all(np.array_equal(c.inliers[k], cr.inliers[k]) for k in range(min(c.n, cr.n)))
        mismatched_frames += int(not same)  # oracle as the checker
    cpu_s = time.perf_counter() - t0
    cpu_value = ns * N_PTS / cpu_s / 1e6

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic (seeded analytic ray casting, SURVEY.md 8d C2)",
        "config": {"workload": "C2: batch of %d 640x480 organized depth frames per GPU, normals+planes+clusters+region records, "
                               "planar seeds %s, frame-wise sharding" % (F, args.seed_mode),
                   "frames_per_gpu": F, "rows": ROWS, "cols": COLS, "points_per_step": total_pts,
                   "l2": "inputs larger than L2 (%.0f MB of points per GPU per step)" % (F * N_PTS * 12 / 1e6),
                   "workspaces_in_flight": D, "numa": numa},
        "value_spread_ms_per_step": dev_spread,
        "batch_latency_ms": batch_latency_ms,
        "p50_frame_ms": batch_latency_ms,
        "p50_frame_ms_definition": "all frames of the batch submitted up front; every frame's result completes with its "
                                   "batch, so completion - submit is the synchronous batch latency for each of them",
        "latency_ms_single_frame": statistics.median(lat),
        "stage_ms": stage_ms, "stage_ms_in_timed_region_overlapped": stage_ms_overlapped,
        "regions_per_frame": {"planar": float(stats[:, 0].mean()), "cluster": float(stats[:, 1].mean())},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(F * N_PTS * 12) * world,
                "d2h_bytes_per_step": e2e_d2h, "ms_per_step": e2e_ms_step, "spread_ms_per_step": e2e_spread,
                "outputs": "label grid + every frame's PlanarRegion/ClusterRegion (records, inlier / boundary / "
                           "discontinuity index lists, convex hulls) to pinned host memory",
                "regions_per_step": {"planar": regions_total[0], "cluster": regions_total[1]}, "ok": e2e_ok,
                "copy_gbs_achieved": e2e_copy_gbs},
        "e2e_variants": variants,
        "host_copy_probe": {"gbs_both_directions_all_ranks": host_copy_gbs,
                            "what": "4 x 256 MB pinned H2D + 4 x 256 MB D2H concurrently per rank, torch copies"},
        "gpu_launches": int(launches),
        "roofline": roofline, "roofline_all": roof,
        "cpu_baseline": {"value": cpu_value, "unit": UNIT, "cores": 1, "kind": cpu_kind(),
                         "sample": "%d frames of the same batch, single-threaded (the reference's native mode), %.1f s" % (ns, cpu_s)},
        "parity": {"frames_checked_vs_oracle": ns, "frames_with_label_or_region_mismatch": mismatched_frames,
                   "checked": "labels, region counts, every region's inlier index list"},
        "clocks": clocks, "host_cores": os.cpu_count(),
    }
    print(json.dumps(line), flush=True)
    for b in batches:
        b.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
