"""CPU: the exact-ties kd-tree build (velociraptor-stf_b200/csrc/qsel.cuh, VRGPU_TREE_TIES=1) run as a serial emulation --
the SAME per-thread source the CUDA kernels call, compiled by g++ -- against the oracle's BuildNodes / MedianPos
restatement (itself bit-identical to the compiled reference on tied inputs: test_oracle_vs_ref.py).
Bar: the identical arrangement (tree position -> particle) and the identical set of nodes, for every thread order."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
QS_SMALL = 2048


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("qsel") / "libqsel_emul.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-Wall", "-I",
                    os.path.join(ROOT, "velociraptor-stf_b200", "csrc"), os.path.join(ROOT, "tests", "host", "qsel_emul.cpp"),
                    "-o", so], check=True)
    lib = C.CDLL(so)
    pll, pd = C.POINTER(C.c_longlong), C.POINTER(C.c_double)
    lib.qsel_emul.restype = C.c_longlong
    lib.qsel_emul.argtypes = [pd, C.c_longlong, pll, C.c_longlong, C.c_int, C.c_int, C.c_uint, C.c_uint, pll, pll, C.c_longlong, pll]

    def run(pos, goff, bucket=8, order_mode=0, seed=1, small_limit=QS_SMALL):
        pos = np.ascontiguousarray(pos, dtype=np.float64)
        goff = np.ascontiguousarray(goff, dtype=np.int64)
        n = pos.shape[0]
        perm = np.zeros(n, dtype=np.int64)
        cap = 4 * n // max(1, bucket) + 64 + len(goff)
        nodes = np.zeros((cap, 2), dtype=np.int64)
        stats = np.zeros(6, dtype=np.int64)
        nn = lib.qsel_emul(pos.ctypes.data_as(pd), n, goff.ctypes.data_as(pll), len(goff) - 1, bucket, order_mode, seed, small_limit,
                           perm.ctypes.data_as(pll), nodes.ctypes.data_as(pll), cap, stats.ctypes.data_as(pll))
        assert nn >= 0, f"emulation failed with {nn}"
        assert nn <= cap
        return perm, nodes[:nn], stats

    return run


def _check(emul, oracle, pos, goff, bucket=8, **kw):
    perm, nodes, stats = emul(pos, goff, bucket=bucket, **kw)
    got_nodes = set(map(tuple, nodes.tolist()))
    want_nodes = set()
    for g in range(len(goff) - 1):
        s, e = int(goff[g]), int(goff[g + 1])
        operm, onodes = oracle.tree_order(pos[s:e], bucket)
        assert np.array_equal(perm[s:e], operm + s), f"group {g}: arrangement differs from MedianPos"
        want_nodes |= {(a + s, b + s) for a, b in onodes.tolist()}
    assert got_nodes == want_nodes
    return stats


def _quantised(n, rng, sigma=0.5):
    """float32 positions of a concentrated clump: exactly tied coordinates along every axis at n >~ 10^4"""
    return (50.0 + rng.normal(size=(n, 3)) * sigma).astype(np.float32).astype(np.float64)


@pytest.mark.parametrize("order_mode", [0, 1, 2])
def test_one_group_with_ties(emul, oracle, order_mode):
    rng = np.random.default_rng(7)
    pos = _quantised(30000, rng, 0.05)
    assert min(len(np.unique(pos[:, d])) for d in range(3)) < 30000
    stats = _check(emul, oracle, pos, [0, 30000], order_mode=order_mode, seed=5)
    assert stats[1] > 0 and stats[3] > 0     # both the parallel rounds and the serial selection ran


@pytest.mark.parametrize("small_limit", [0, 5, 64, 100000])
def test_every_size_takes_both_paths(emul, oracle, small_limit):
    """small_limit 0: every selection by parallel rounds (down to 2-particle ranges); 100000: everything serial."""
    rng = np.random.default_rng(8)
    pos = _quantised(6000, rng, 0.01)
    stats = _check(emul, oracle, pos, [0, 6000], order_mode=2, seed=3, small_limit=small_limit)
    if small_limit == 0:
        assert stats[3] == 0
    if small_limit == 100000:
        assert stats[1] == 0


def test_several_groups_and_buckets(emul, oracle):
    rng = np.random.default_rng(9)
    sizes = [9000, 5000, 2500, 2049, 2048, 151, 17, 9, 8, 3]
    goff = np.concatenate([[0], np.cumsum(sizes)])
    pos = np.concatenate([_quantised(s, rng, 0.02) + k for k, s in enumerate(sizes)])
    for bucket in (8, 16, 1):
        _check(emul, oracle, pos, goff, bucket=bucket, order_mode=2, seed=bucket)


def test_degenerate_inputs(emul, oracle):
    rng = np.random.default_rng(10)
    n = 5000
    same = np.full((n, 3), 3.25)                                               # every particle at one point
    plane = np.stack([np.full(n, 1.0), rng.integers(0, 5, n).astype(float), rng.random(n)], axis=1)
    lattice = rng.integers(0, 7, size=(n, 3)).astype(float)                    # at most 343 distinct points
    for pos in (same, plane, lattice):
        for small_limit in (0, QS_SMALL):
            _check(emul, oracle, pos, [0, n], order_mode=2, seed=1, small_limit=small_limit)


def test_distinct_coordinates(emul, oracle):
    rng = np.random.default_rng(11)
    pos = rng.random((20000, 3)) * 100.0
    stats = _check(emul, oracle, pos, [0, 12000, 20000], order_mode=2, seed=2)
    assert stats[4] < 200                    # rounds per level stay logarithmic
    assert stats[5] < 6 * 20000 * 4          # work items: a few n per level with nodes above QS_SMALL (4 such levels), not rounds x n


def test_random_small_inputs(emul, oracle):
    rng = np.random.default_rng(12)
    for trial in range(200):
        n = int(rng.integers(1, 400))
        kind = trial % 4
        if kind == 0:
            pos = rng.random((n, 3))
        elif kind == 1:
            pos = rng.integers(0, 3, size=(n, 3)).astype(float)
        elif kind == 2:
            pos = np.round(rng.normal(size=(n, 3)) * 2.0) / 2.0
        else:
            pos = _quantised(n, rng, 1e-5)
        cut = int(rng.integers(0, n + 1))
        goff = [0, cut, n] if 0 < cut < n else [0, n]
        _check(emul, oracle, pos, goff, bucket=int(rng.integers(1, 10)), order_mode=trial % 3, seed=trial,
               small_limit=int(rng.choice([0, 3, 50, QS_SMALL])))
