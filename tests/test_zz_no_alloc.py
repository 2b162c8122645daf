"""pcseg_seeds_average_normals_noalloc (include/pcseg_b200.h): the average-normals seed finder with
its scratch in a static device array -- no counterpart in the reference, added for callers that
cannot call cudaMalloc.  Checked through the C-ABI against pcseg_seeds_average_normals: the seed
lists are index work, so the comparison is bit-exact (np.array_equal), on calls in a row that
share the static scratch.  Collected last (test_zz_*) on purpose."""
import numpy as np
import pytest

import pcseg
import synth


@pytest.fixture()
def ctx():
    c = pcseg.Context(0)
    yield c
    c.close()


def _case(ctx, rows, cols, seed):
    xyz = synth.depth_frame(rows, cols, seed=seed)
    nrm = ctx.normals_organized(xyz, rows, cols, params=dict(min_neighbor_distance=0.05))
    return xyz, np.ascontiguousarray(nrm, np.float32)


def _noalloc(ctx, torch, nrm, rows, cols, nbhd, min_valid, min_len, check_free=False):
    """Caller-owned device buffers (torch is only the caller's allocator here); the entry point
    itself must leave the device's free memory untouched."""
    d_nrm = torch.from_numpy(nrm.reshape(-1)).cuda()
    cap = rows * cols
    d_seeds = torch.full((cap,), -7, dtype=torch.int32, device="cuda")
    d_n = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    ctx.seeds_average_normals_noalloc(d_nrm.data_ptr(), rows, cols, d_seeds.data_ptr(), cap, d_n.data_ptr(),
                                      nbhd, min_valid, min_len)
    free1 = torch.cuda.mem_get_info()[0]
    if check_free:
        assert free1 == free0, "the no-alloc entry point changed the device's free memory"
    n = int(d_n.cpu()[0])
    assert 0 <= n <= cap
    out = d_seeds.cpu().numpy()
    assert np.all(out[n:] == -7)  # nothing written behind the seeds
    return out[:n].copy()


@pytest.mark.gpu
def test_noalloc_matches_existing_entry_point_on_calls_in_a_row(ctx):
    import torch
    # a large frame first, then a smaller one with another window: the second call finds the first
    # call's flags in the shared array and must not see them; then the first case again
    cases = [(240, 320, 7, 11), (96, 128, 5, 12), (240, 320, 7, 11), (64, 64, 3, 13)]
    for k, (rows, cols, nbhd, seed) in enumerate(cases):
        xyz, nrm = _case(ctx, rows, cols, seed)
        want = ctx.seeds_average_normals(xyz, nrm, rows, cols, nbhd, 8, 0.999)
        got = _noalloc(ctx, torch, nrm, rows, cols, nbhd, 8, 0.999, check_free=(k > 0))
        assert len(want) > 50
        assert got.dtype == want.dtype and np.array_equal(got, want), (rows, cols, nbhd)


@pytest.mark.gpu
def test_noalloc_two_contexts_share_the_scratch(ctx):
    """The array belongs to the device, not to a context: a second context gets the same results."""
    import torch
    other = pcseg.Context(0)
    try:
        for c, (rows, cols, nbhd, seed) in ((ctx, (120, 160, 5, 21)), (other, (96, 128, 7, 22)), (ctx, (96, 128, 7, 22))):
            xyz, nrm = _case(c, rows, cols, seed)
            want = c.seeds_average_normals(xyz, nrm, rows, cols, nbhd, 8, 0.999)
            got = _noalloc(c, torch, nrm, rows, cols, nbhd, 8, 0.999)
            assert np.array_equal(got, want)
    finally:
        other.close()


@pytest.mark.gpu
def test_noalloc_limits(ctx):
    import torch
    L = ctx.L
    d = torch.zeros(16, dtype=torch.float32, device="cuda")
    di = torch.zeros(16, dtype=torch.int32, device="cuda")
    # beyond the static scratch: rejected before anything is read
    rows, cols = 1024, 1025
    with pytest.raises(pcseg.PcsegError) as e:
        ctx.seeds_average_normals_noalloc(d.data_ptr(), rows, cols, di.data_ptr(), rows * cols, di.data_ptr())
    assert e.value.status == 5 and "PCSEG_NOALLOC_MAX_CELLS" in str(e.value)
    # cap below the number of window positions
    with pytest.raises(pcseg.PcsegError) as e:
        ctx.seeds_average_normals_noalloc(d.data_ptr(), 96, 128, di.data_ptr(), 16, di.data_ptr())
    assert e.value.status == 5
    # host pointers would need a staging allocation
    h = np.zeros(96 * 128 * 3, np.float32)
    with pytest.raises(pcseg.PcsegError) as e:
        ctx.seeds_average_normals_noalloc(h.ctypes.data, 96, 128, di.data_ptr(), 96 * 128, di.data_ptr())
    assert e.value.status == 6
    # fewer rows than the window: no seeds (ps/segmentation.h:150-152), *n_out = 0
    nrm = torch.zeros(4 * 64 * 3, dtype=torch.float32, device="cuda")
    seeds = torch.zeros(4 * 64, dtype=torch.int32, device="cuda")
    n = torch.full((1,), 9, dtype=torch.int32, device="cuda")
    ctx.seeds_average_normals_noalloc(nrm.data_ptr(), 4, 64, seeds.data_ptr(), 4 * 64, n.data_ptr(), 5)
    assert int(n.cpu()[0]) == 0
    assert L is not None
