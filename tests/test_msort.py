"""The sort of SlidingMeanShift (ps/mean_shift_segmentation.h:276-280).  Its permutation is the
standard library's (quirk Q7), so the checker is the real std::sort of this box (oracle.sort_order).
CPU: the level-parallel restatement (oracle/msort_model.py) reproduces it; GPU: the device sort
(csrc/msort.cuh through pcseg_mean_shift_sort_order) reproduces it, bit-exact permutations."""
import numpy as np
import pytest

import oracle
import msort_model


def cases(big):
    rng = np.random.default_rng(0)
    out = []
    for n in [0, 1, 5, 16, 17, 18, 33, 100, 1000, 5000] + ([50000, 300000, 2000000] if big else [20000]):
        out.append(("window counts %d" % n, rng.integers(50, 122, n)))
        out.append(("3 keys %d" % n, rng.integers(0, 3, n)))
    k = np.concatenate([np.full(30000, 60), rng.integers(50, 122, 3000)])
    rng.shuffle(k)
    out.append(("one dominant key", k))
    k = np.sort(rng.integers(50, 122, 20000))
    out.append(("ascending", k))
    out.append(("descending", k[::-1].copy()))
    out.append(("all equal", np.full(4097, 77)))
    out.append(("full byte range", rng.integers(0, 256, 70000)))
    return out


def test_level_parallel_restatement_matches_std_sort():
    for name, k in cases(False):
        assert np.array_equal(msort_model.replay(k), oracle.sort_order(k.astype(np.float64), 0)), name


@pytest.fixture(scope="module")
def ctx():
    import pcseg
    c = pcseg.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_cuda_sort_matches_std_sort(ctx):
    rng = np.random.default_rng(3)
    for name, k in cases(True):
        # sprinkle invalid seeds (intensity -1) between the valid ones: they are dropped first
        full = np.full(k.size + k.size // 3 + 2, -1.0, np.float32)
        pos = np.sort(rng.choice(full.size, k.size, replace=False))
        full[pos] = k
        for mode in (0, 1):
            got = ctx.mean_shift_sort_order(full, mode)
            want = pos[oracle.sort_order(k.astype(np.float64), mode)]
            assert np.array_equal(got, want), (name, mode)


@pytest.mark.gpu
def test_cuda_sort_rejects_non_count_keys(ctx):
    import pcseg
    with pytest.raises(pcseg.PcsegError):
        ctx.mean_shift_sort_order(np.array([50.5, 60.0], np.float32), 0)
