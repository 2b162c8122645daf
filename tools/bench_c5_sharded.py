"""C5 on N GPUs (torchrun): SlidingMeanShift of the 10 M-point cloud with the seed cells
sharded over the ranks, cloud + labels replicated, mode records all-gathered with NCCL
(parallel.py), sort + growth once on rank 0.  Checks the gathered records and the final labels
against the unsharded single-GPU result and prints one JSON object on rank 0.
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bench_c5_sharded.py"""
import ctypes as C
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200"))
import numpy as np
import torch
import torch.distributed as dist
import parallel, pcseg, synth

rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rows, cols, iters = 2500, 4000, 5
n = rows * cols
xyz = synth.c5_full(rows, cols)                     # same seeded cloud on every rank (replicated)
d_xyz = torch.from_numpy(xyz).cuda()
d_lab = torch.full((n,), -1, dtype=torch.int32, device="cuda")
ctx = pcseg.Context(local)
L = ctx.L


def modes_fn(b, e, out):
    ctx._check(L.pcseg_mean_shift_modes_range(ctx.h, C.c_void_p(d_xyz.data_ptr()), rows, cols,
                                              C.c_void_p(d_lab.data_ptr()), iters, b, e, C.c_void_p(out.data_ptr())))


def iteration_phase():
    return parallel.sharded_mean_shift_modes(modes_fn, n, rank, world, device=d_xyz.device)

for _ in range(2):
    iteration_phase()
times = []
for _ in range(5):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    modes = iteration_phase()
    e1.record(); e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    times.append(float(t.item()))
ms = float(np.median(times))
t0 = time.perf_counter()
labels, regs = parallel.sharded_mean_shift(ctx, d_xyz, rows, cols, d_lab, rank, world, iterations=iters)
torch.cuda.synchronize()
total_ms = (time.perf_counter() - t0) * 1e3
if rank == 0:
    full = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    ctx._check(L.pcseg_mean_shift_modes(ctx.h, C.c_void_p(d_xyz.data_ptr()), rows, cols, C.c_void_p(d_lab.data_ptr()),
                                        iters, C.c_void_p(full.data_ptr())))
    same_modes = bool(torch.equal(full.view(torch.int32), modes.view(torch.int32)))
    lab1, regs1 = ctx.mean_shift(xyz, rows, cols, np.full(n, -1, np.int32), iterations=iters)
    same_labels = bool(np.array_equal(lab1, labels)) and regs1.n == regs.n
    seeds = int((~np.isnan(xyz[:, 0])).sum())
    print(json.dumps({"config": "C5 mean shift, 2500x4000 = 10 M points, %d iterations" % iters, "n_gpus": world,
                      "iteration_phase_ms_max_over_ranks": ms, "Gseeds_per_s": seeds / ms / 1e6,
                      "allgather_bytes_per_rank": int(-(-n // world) * 32), "sliding_mean_shift_total_ms": total_ms,
                      "regions": int(regs.n), "sharded_records_bit_identical_to_unsharded": same_modes,
                      "labels_identical_to_single_gpu": same_labels}))
if world > 1:
    dist.destroy_process_group()
