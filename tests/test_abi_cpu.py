"""CPU checks of the drop-in boundary: libibl.so loads, exports exactly what include/ibl.h declares, and the product
path fails loudly (never falls back) when there is no GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "ibl.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ibl_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from instance_based_loc_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/ibl.h but not exported"
    # and the ctypes prototype table covers the header
    assert sorted(_lib.EXPORTS) == syms


def test_version_and_error_string():
    from instance_based_loc_b200 import _lib
    L = _lib.lib()
    assert L.ibl_version() >= 100
    assert L.ibl_prof_num_tags() > 10 and L.ibl_launch_count() >= 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from instance_based_loc_b200 import kernels
    from instance_based_loc_b200._lib import IblError
    from instance_based_loc_b200.o3d.pointcloud import PointCloud
    with pytest.raises(IblError):
        kernels.to_dev(np.zeros((4, 3)))
    with pytest.raises(IblError):
        PointCloud(np.zeros((4, 3)))


def test_product_never_imports_oracle():
    """the oracle is test infrastructure: no module of the product package may import it."""
    pkg = os.path.join(ROOT, "instance_based_loc_b200")
    bad = []
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                txt = open(os.path.join(d, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M):
                    bad.append(os.path.join(d, f))
    assert not bad, bad


def test_simvolume_bookkeeping_matches_oracle():
    """Host logic of SimVolume.get_top_indices_from_subvolumes (vectorised first-occurrence / per-length selection) against
    the oracle's literal restatement of similarity_volume.py:226-270, fed with the oracle's own top-k entries."""
    import numpy as np
    from oracle import ref_port
    from instance_based_loc_b200.utils.similarity_volume import assignments_from_topk
    rng = np.random.default_rng(5)
    checked = 0
    for trial in range(80):
        n = int(rng.integers(1, 8))
        m = int(rng.integers(max(n, 3), 24))
        S = rng.uniform(-0.3, 1, (n, m)).astype(np.float32)
        if trial % 3 == 0:
            S = np.round(S, 1)                                  # many exact ties after the fp16 roundings
        chosen, subs = ref_port.simvolume_subvolumes(S)
        k = 4 * n * 4
        if min(v.size for v in subs) < k:
            continue
        top = [ref_port.topk_first_index(v, k) for v in subs]
        flat = np.stack([np.asarray(t[0], np.int64) for t in top])
        val = np.stack([np.asarray(t[1], np.float32) for t in top])
        got = assignments_from_topk(flat, val, chosen, m, len(chosen[0]), n)
        assert got == ref_port.simvolume_assignments(S, 4), (trial, n, m)
        checked += 1
    assert checked > 40
