"""CPU-side checks of the C-ABI library: it loads and exports every symbol that
include/pcseg_b200.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest

import pcseg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "pcseg_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pcseg_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = pcseg.lib()
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(L, n), n
    assert sorted(pcseg.EXPORTS) == names


def test_defaults_match_reference():
    p = pcseg.normals_params()
    assert (p.min_neighbor_distance, p.max_neighbor_distance) == (ctypes.c_float(0.1).value, 1.0)
    assert (p.include_diagonal_neighbors, p.min_num_support_neighbors) == (1, 4)
    c = pcseg.planar_config()
    assert c.min_region_inliers == 5 and c.plane_model_reestimation_period == 30
    assert c.max_plane_distance == ctypes.c_float(0.05).value
    k = pcseg.cluster_config()
    assert (k.min_region_inliers, k.squared_distance_threshold, k.half_search_window) == (7, 1.0, 1)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pcseg.PcsegError):
        pcseg.Context(0)


def test_product_does_not_reference_oracle():
    pkg = os.path.join(ROOT, "x-edr-pointcloud-segmentation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in txt and "import oracle" not in txt, f
