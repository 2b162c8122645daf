"""TEST INFRASTRUCTURE ONLY: level-by-level restatement of libstdc++'s std::sort (GCC 13
bits/stl_algo.h:1918-1951) as the device kernels in csrc/msort.cuh compute it -- every recursion
level of __introsort_loop as one data-parallel pass stated by ranks.  tests/test_msort.py pins it
against the real std::sort (oracle.sort_order) on this box."""
import numpy as np


def replay(keys):
    """keys: integer array.  Returns order[k] = input position of the element std::sort leaves at k."""
    n = len(keys)
    key = np.array(keys, np.int64)
    idx = np.arange(n)
    if n == 0:
        return idx
    depth_limit = 2 * (int(n).bit_length() - 1)   # 2 * std::__lg(n)
    segs = [(0, n)]
    level = 0
    while segs:
        nxt = []
        for lo, hi in segs:
            if hi - lo <= 16:                       # _S_threshold
                continue
            if level >= depth_limit:
                raise RuntimeError("std::sort would fall back to heapsort here")
            # __move_median_to_first(first, first + 1, mid, last - 1), :84-108
            a, b, c = lo + 1, lo + (hi - lo) // 2, hi - 1
            ka, kb, kc = key[a], key[b], key[c]
            if ka < kb:
                m = b if kb < kc else (c if ka < kc else a)
            elif ka < kc:
                m = a
            elif kb < kc:
                m = c
            else:
                m = b
            key[[lo, m]] = key[[m, lo]]
            idx[[lo, m]] = idx[[m, lo]]
            p = key[lo]
            # __unguarded_partition(first + 1, last, first), :1871-1889, by ranks
            rng = np.arange(lo + 1, hi)
            lpos = rng[key[rng] >= p]               # stoppers of the left scan, ascending
            rpos = rng[key[rng] <= p][::-1]         # stoppers of the right scan, descending
            m_ = min(len(lpos), len(rpos))
            s = int(np.sum(lpos[:m_] < rpos[:m_]))  # pairs that are swapped
            cut = lpos[s] if s < len(lpos) else 1 << 60
            if s >= 1:
                cut = min(cut, rpos[s - 1])
            cut = int(cut)
            i, j = lpos[:s], rpos[:s]
            key[i], key[j] = key[j].copy(), key[i].copy()
            idx[i], idx[j] = idx[j].copy(), idx[i].copy()
            nxt.append((cut, hi))
            nxt.append((lo, cut))
        segs = nxt
        level += 1
    # __final_insertion_sort: the stable sort of the arrangement the loop left
    return idx[np.argsort(key, kind="stable")]
