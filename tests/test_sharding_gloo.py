"""Host-side logic of the multi-GPU paths, exercised with world_size-2 gloo
process groups on the CPU (no GPU needed): frame sharding and the all-gather of
the sharded mean-shift mode records (parallel.py).  The per-shard records come
from the CPU oracle here (the checker); on GPUs they come from
pcseg_mean_shift_modes_range (tests/test_gpu_parity.py covers that kernel)."""
import os
import socket

import numpy as np
import pytest

import oracle
import parallel
import synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _full_records(xyz, rows, cols, iterations):
    """Oracle mode records in the product's per-cell layout (8 floats, is_seed last)."""
    n = rows * cols
    lab = np.full(n, -1, np.int32)
    _, _, modes = oracle.mean_shift(xyz, rows, cols, lab, iterations=iterations)
    rec = np.zeros((n, 8), np.float32)
    cells = modes[:, 7].astype(np.int64)
    rec[cells, :7] = modes[:, :7]
    rec[cells, 7] = 1.0
    return rec


def _worker(rank, world, port, rows, cols, iterations, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        xyz = synth.c5(rows, cols, seed=4234, n_blobs=8, pitch=0.12)
        rec = _full_records(xyz, rows, cols, iterations)
        calls = []

        def modes_fn(b, e, out):
            calls.append((b, e))
            out.copy_(torch.from_numpy(rec[b:e]))

        full = parallel.sharded_mean_shift_modes(modes_fn, rows * cols, rank, world)
        np.save(os.path.join(out_dir, "full_%d.npy" % rank), full.numpy())
        np.save(os.path.join(out_dir, "calls_%d.npy" % rank), np.array(calls))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("rows,cols", [(90, 120), (45, 67)])  # 45*67 = 3015 cells: uneven shards
def test_sharded_modes_all_gather_world2(tmp_path, rows, cols):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), rows, cols, 3, str(tmp_path)), nprocs=world, join=True)
    xyz = synth.c5(rows, cols, seed=4234, n_blobs=8, pitch=0.12)
    want = _full_records(xyz, rows, cols, 3)
    ranges = parallel.cell_ranges(rows * cols, world)
    for r in range(world):
        got = np.load(tmp_path / ("full_%d.npy" % r))
        assert got.shape == want.shape
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))  # bit-exact gather
        calls = np.load(tmp_path / ("calls_%d.npy" % r))
        assert [tuple(c) for c in calls] == [ranges[r]]


def test_cell_ranges_cover_and_are_disjoint():
    for n in (1, 7, 100, 3015, 10_000_000):
        for w in (1, 2, 3, 4, 8):
            rs = parallel.cell_ranges(n, w)
            assert len(rs) == w and rs[0][0] == 0 and rs[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
            assert all(b <= e for b, e in rs)
            per = -(-n // w)
            assert all(e - b == per for b, e in rs[:-1] if e < n)


def test_frame_shard_round_robin():
    for g in (1, 2, 4, 8):
        owned = [parallel.frame_shard(256, r, g) for r in range(g)]
        assert sorted(sum(owned, [])) == list(range(256))
        assert all(i % g == r for r in range(g) for i in owned[r])
