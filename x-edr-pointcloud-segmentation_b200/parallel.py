"""Multi-GPU plumbing of the hot path (one process per GPU, torch.distributed).

Only two things on this path shard:

* batches of independent frames: frame i -> rank i mod G, no collective
  (`frame_shard`); results stay per rank / are gathered by the caller;
* the SlidingMeanShift ITERATION phase of one large cloud
  (ps/mean_shift_segmentation.h:232-260): seeds are mutually independent while
  the labels are read-only, so every rank computes the mode records of one
  contiguous linear-index range of seed cells on a replicated cloud
  (`pcseg_mean_shift_modes_range`) and the records are all-gathered (NCCL over
  NVLink on GPUs, gloo in the CPU tests); the sequential sort + growth phase
  then runs once (`pcseg_mean_shift_from_modes`).

Single-frame region growing does not shard (one FIFO, shared labels):
replicas only.  torch is plumbing here (process group, device buffers).
"""
import numpy as np

MODE_FLOATS = 8  # ModeRecord: mode xyz, index row, index col, intensity, valid, is_seed


def frame_shard(n_frames, rank, world):
    """Frame ids owned by `rank`: frame i -> rank i mod world."""
    return list(range(rank, n_frames, world))


def cell_ranges(n_cells, world):
    """Contiguous [begin, end) seed-cell ranges, equal length except the tail.
    All shards but the last have ceil(n/world) cells so that the gather buffer
    is a plain concatenation."""
    per = -(-n_cells // world)
    return [(min(r * per, n_cells), min((r + 1) * per, n_cells)) for r in range(world)]


def sharded_mean_shift_modes(modes_fn, n_cells, rank, world, group=None, device=None):
    """All-gathers the mode records of the per-rank seed-cell ranges.

    modes_fn(cell_begin, cell_end, out) fills `out`, a float32 torch tensor
    [cell_end - cell_begin, 8] living on `device`, with this rank's records
    (on GPUs: pcseg_mean_shift_modes_range writing straight into out.data_ptr()).
    Returns the full [n_cells, 8] tensor on every rank."""
    import torch
    import torch.distributed as dist
    ranges = cell_ranges(n_cells, world)
    per = -(-n_cells // world)
    b, e = ranges[rank]
    full = torch.zeros((world * per, MODE_FLOATS), dtype=torch.float32, device=device)
    mine = full[rank * per: rank * per + per]
    if device is not None and torch.device(device).type == "cuda":
        # the zero fill runs on torch's stream, the kernel on the library's own stream (which the
        # call synchronises before it returns): order them explicitly
        torch.cuda.current_stream().synchronize()
    if e > b:
        modes_fn(b, e, mine[: e - b])
    if world > 1:
        # in-place all-gather: every rank contributes its own slice of `full`
        dist.all_gather_into_tensor(full, mine.clone(), group=group)
    return full[:n_cells]


def sharded_mean_shift(ctx, xyz_dev, rows, cols, labels_dev, rank, world, cfg=None, iterations=5,
                       id_offset=0, sort_mode=0, group=None, grow_on_all_ranks=False, labels_out=None,
                       labels_on_device=False):
    """SlidingMeanShift of one large cloud with the seed cells sharded over the
    ranks.  xyz_dev / labels_dev: torch CUDA tensors (replicated on every rank).
    Returns (labels numpy, regions) on rank 0 (on every rank if
    grow_on_all_ranks), else (None, None).  labels_out: a pinned host tensor to receive the grid;
    labels_on_device: return the CUDA tensor instead of a host copy."""
    import ctypes as C
    import torch
    import pcseg
    n = rows * cols
    L = ctx.L

    def modes_fn(b, e, out):
        st = L.pcseg_mean_shift_modes_range(ctx.h, C.c_void_p(xyz_dev.data_ptr()), rows, cols,
                                            C.c_void_p(labels_dev.data_ptr()), iterations, b, e,
                                            C.c_void_p(out.data_ptr()))
        ctx._check(st)

    modes = sharded_mean_shift_modes(modes_fn, n, rank, world, group=group, device=xyz_dev.device)
    if rank != 0 and not grow_on_all_ranks:
        return None, None
    modes = modes.contiguous()
    labels = labels_dev.clone()
    torch.cuda.current_stream().synchronize()   # clone / gather done before the library's stream reads them
    c = cfg if isinstance(cfg, pcseg.ClusterConfig) else pcseg.cluster_config(**(cfg or {}))
    h = C.c_void_p()
    ctx._check(L.pcseg_mean_shift_from_modes(ctx.h, C.c_void_p(xyz_dev.data_ptr()), rows, cols,
                                             C.c_void_p(modes.data_ptr()), C.byref(c), id_offset, sort_mode,
                                             C.c_void_p(labels.data_ptr()), C.byref(h)))
    regs = pcseg._collect(L, h, "meanshift")
    L.pcseg_regions_free(h)
    torch.cuda.synchronize()
    if labels_out is not None:   # caller's (pinned) host tensor
        labels_out.copy_(labels)
        return labels_out.numpy(), regs
    if labels_on_device:         # the label grid stays in HBM (a torch tensor)
        return labels, regs
    return labels.cpu().numpy(), regs
