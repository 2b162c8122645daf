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


def _cpu_pipeline(xyz):
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
    out_normals = torch.empty((F, N_PTS, 3), dtype=torch.float32).pin_memory()
    out_labels = torch.empty((F, N_PTS), dtype=torch.int32).pin_memory()

    ctx = pcseg.Context(local_rank)
    D = args.depth
    batches = [pcseg.Batch(ctx, F, ROWS, COLS) for _ in range(D)]
    batch = batches[0]
    out_normals, out_labels = [out_normals], [out_labels]
    try:  # one pinned output set per workspace; a host that cannot pin that much runs with fewer in flight
        for _ in range(D - 1):
            out_normals.append(torch.empty((F, N_PTS, 3), dtype=torch.float32).pin_memory())
            out_labels.append(torch.empty((F, N_PTS), dtype=torch.int32).pin_memory())
    except RuntimeError:
        D = min(len(out_normals), len(out_labels))
        out_normals, out_labels = out_normals[:D], out_labels[:D]
        for b in batches[D:]:
            b.close()
        batches = batches[:D]
    mode = pcseg.SEEDS_ALL_DESCENDING if args.seed_mode == "alldesc" else pcseg.SEEDS_PLANE_SUPPORT
    params = pcseg.pipeline_params(seed_mode=mode, run_cluster=True)
    stream = torch.cuda.ExternalStream(batch.stream)

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

    stage_acc = {"normals": 0.0, "seeds": 0.0, "planar": 0.0, "cluster": 0.0}

    def run_steps(steps, submit, collect=False):
        """`steps` passes of the pipeline, step i on workspace i % D; a workspace is waited for
        (and its stage timers read) before it is reused, and all are drained at the end."""
        busy = [False] * D
        for i in range(steps):
            j = i % D
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

    def timed(submit, steps, collect=False):
        """K steps bracketed by barrier+sync; device time between two CUDA events on the first
        workspace's stream (idle at both records, so they bracket the whole region); also the
        host wall clock."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        run_steps(steps, submit, collect)
        e1.record(stream)
        e1.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)), max_over_ranks(wall)

    # ---- device-resident arm (value) ----
    submit_dev = lambda j: batches[j].submit(dev_xyz.data_ptr(), params)
    run_steps(max(args.warmup, D), submit_dev)   # every workspace allocates its buffers on first use
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    dev_ms, dev_wall = timed(submit_dev, args.steps, collect=True)
    launches = ctx.launches - l0
    clocks = sampler.stop() if rank == 0 else None
    stats = batch.frame_stats()
    ms_per_step = dev_ms / args.steps
    total_pts = F * N_PTS * world
    value = total_pts / (ms_per_step * 1e-3) / 1e6

    # ---- end-to-end arm (host buffers through the C ABI): every step copies its 256 frames in
    # from pinned host memory and its result out.  The result of the path is the label grid (the
    # regions' inlier sets; PlanarRegion/ClusterRegion come out, north_star); the normals are an
    # intermediate that a caller may also ask for -- measured as a second variant. ----
    host_np = host.numpy()
    on_np = [t.numpy() for t in out_normals]
    ol_np = [t.numpy() for t in out_labels]
    submit_e2e = lambda j: batches[j].submit(host_np, params, labels_out=ol_np[j])
    run_steps(max(2, D), submit_e2e)             # staging buffers of every workspace exist before timing
    e2e_ms, e2e_wall = timed(submit_e2e, args.steps)
    e2e_ms_step = max(e2e_ms, e2e_wall) / args.steps
    e2e_value = total_pts / (e2e_ms_step * 1e-3) / 1e6
    submit_e2e_n = lambda j: batches[j].submit(host_np, params, normals_out=on_np[j], labels_out=ol_np[j])
    run_steps(max(2, D), submit_e2e_n)
    e2en_ms, e2en_wall = timed(submit_e2e_n, args.steps)
    e2en_ms_step = max(e2en_ms, e2en_wall) / args.steps
    e2en_value = total_pts / (e2en_ms_step * 1e-3) / 1e6
    e2e_ok = int(stats[:, 5].max()) == 0
    ol_np = ol_np[0]

    # ---- one batch alone, synchronous: its latency (all frames submitted up front) ----
    sync_ms = []
    for it in range(3):
        t0 = time.perf_counter()
        batch.run(dev_xyz.data_ptr(), params)
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
    roof = {}
    for k in ("normals", "planar") if fused else ("normals", "planar", "cluster"):
        ach = alg_bytes[k] / (stage_ms[k] * 1e-3) / 1e9 if stage_ms[k] > 0 else 0.0
        roof[k] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                   "traffic": None, "kernel_ms": stage_ms[k], "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)"}
    roofline = dict(roof[dom])
    roofline["kernel"] = {"planar": "planar_grow_kernel" + (" (planar + fused cluster growth)" if fused else ""),
                          "normals": "normals_kernel", "cluster": "cluster_grow_kernel"}[dom]
    roofline["note"] = ("kernel_ms: one batch alone (CUDA events on its stream); latency-bound serial FIFO chain "
                        "(one re-fit per 30 inliers) whose top stall is instruction fetch, not bandwidth-bound; "
                        "see DESIGN.md section 8 and profiles/")

    # ---- CPU baseline: oracle port, single thread, bounded sample ----
    ns = min(args.cpu_sample, F)
    t0 = time.perf_counter()
    mismatched_frames = 0
    for i in range(ns):
        lab_cpu, _, _ = _cpu_pipeline(host_np[i])
        mismatched_frames += int(not np.array_equal(lab_cpu, ol_np[i]))  # oracle as the checker
    cpu_s = time.perf_counter() - t0
    cpu_value = ns * N_PTS / cpu_s / 1e6

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic (seeded analytic ray casting, SURVEY.md 8d C2)",
        "config": {"workload": "C2: batch of %d 640x480 organized depth frames per GPU, normals+planes+clusters, "
                               "planar seeds %s, frame-wise sharding" % (F, args.seed_mode),
                   "frames_per_gpu": F, "rows": ROWS, "cols": COLS, "points_per_step": total_pts,
                   "l2": "inputs larger than L2 (%.0f MB of points per GPU per step)" % (F * N_PTS * 12 / 1e6),
                   "workspaces_in_flight": D, "numa": numa},
        "p50_frame_ms": batch_latency_ms, "latency_ms_single_frame": statistics.median(lat),
        "stage_ms": stage_ms, "stage_ms_in_timed_region_overlapped": stage_ms_overlapped,
        "regions_per_frame": {"planar": float(stats[:, 0].mean()), "cluster": float(stats[:, 1].mean())},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(F * N_PTS * 12) * world,
                "d2h_bytes_per_step": int(F * N_PTS * 4) * world, "ms_per_step": e2e_ms_step,
                "outputs": "labels (int32 per point) to pinned host memory", "ok": e2e_ok},
        "e2e_with_normals": {"value": e2en_value, "unit": UNIT, "h2d_bytes_per_step": int(F * N_PTS * 12) * world,
                             "d2h_bytes_per_step": int(F * N_PTS * 16) * world, "ms_per_step": e2en_ms_step,
                             "outputs": "normals + labels to pinned host memory"},
        "gpu_launches": int(launches),
        "roofline": roofline, "roofline_all": roof,
        "cpu_baseline": {"value": cpu_value, "unit": UNIT, "cores": 1, "kind": cpu_kind(),
                         "sample": "%d frames of the same batch, single-threaded (the reference's native mode), %.1f s" % (ns, cpu_s)},
        "parity": {"frames_checked_vs_oracle": ns, "frames_with_label_mismatch": mismatched_frames},
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
