"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/orbx.h
declares, and fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def orbx():
    import __graft_entry__
    __graft_entry__.build()
    from orb_slam3_ros2_b200 import orbx
    return orbx


def _declared():
    text = open(os.path.join(ROOT, "include", "orbx.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(orbx):
    names = _declared()
    assert len(names) >= 25
    l = ctypes.CDLL(orbx.LIB_PATH)
    for n in names:
        assert hasattr(l, n), f"{n} declared in include/orbx.h but not exported"
    assert sorted(orbx.SYMBOLS) == names  # the ctypes table covers the header exactly


def test_host_side_hamming_matches_descriptor_distance(orbx, oracle):
    rng = np.random.default_rng(1)
    for _ in range(50):
        a, b = rng.integers(0, 256, 32, dtype=np.uint8), rng.integers(0, 256, 32, dtype=np.uint8)
        assert orbx.DescriptorDistance(a, b) == oracle.hamming(a, b) == int(np.unpackbits(a ^ b).sum())
    z = np.zeros(32, dtype=np.uint8)
    assert orbx.DescriptorDistance(z, z) == 0 and orbx.DescriptorDistance(z, ~z) == 256


def test_no_cpu_fallback(orbx):
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    with pytest.raises(orbx.OrbxError) as e:
        orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.code == orbx.ORBX_ERR_CUDA
    with pytest.raises(orbx.OrbxError):
        orbx.match_top2(np.zeros((4, 32), np.uint8), np.zeros((8, 32), np.uint8))


def test_shard_ranges_partition_the_database(orbx):
    for n in (10, 2000, 10_000_000, 7):
        for g in (1, 2, 4, 8):
            r = [orbx.shard_range(n, k, g) for k in range(g)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(g - 1))
