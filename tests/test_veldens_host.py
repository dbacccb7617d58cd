"""CPU: the local velocity density (velociraptor-stf_b200/csrc/veldens.cuh behind vrgpu_velocity_density) run
as a serial emulation -- the SAME per-thread source the CUDA kernels call, compiled by g++ -- against the oracle's restatement
of GetVelocityDensity, which is itself bit-identical to the compiled reference (tests/test_oracle_vs_ref.py).
Bar: 1e-12 relative (the only operation that is not reproduced to the bit is the reference's pow(h, 3.0): the device code
cubes with one rounding); in practice almost every density is bit-identical, which the tests also count."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("veldens") / "libveldens_emul.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-Wall", "-I",
                    os.path.join(ROOT, "velociraptor-stf_b200", "csrc"), os.path.join(ROOT, "tests", "host", "veldens_emul.cpp"),
                    "-o", so], check=True)
    lib = C.CDLL(so)
    pd, pi, pll = C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_longlong)
    lib.veldens_emul.restype = C.c_int
    lib.veldens_emul.argtypes = [pd, pd, pi, C.c_longlong, pll, C.c_longlong, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint, pd, pll]

    def run(pos, vel, period=0.0, bsize=16, nsearch=256, nvel=32, approx=1, ptype=None, batch=0, leaf_offsets=None):
        pos = np.ascontiguousarray(pos, dtype=np.float64)
        vel = np.ascontiguousarray(vel, dtype=np.float64)
        n = pos.shape[0]
        tp = None if ptype is None else np.ascontiguousarray(ptype, dtype=np.int32)
        den = np.zeros(n)
        stats = np.zeros(2, dtype=np.int64)
        lo = None if leaf_offsets is None else np.ascontiguousarray(leaf_offsets, dtype=np.int64)
        rc = lib.veldens_emul(pos.ctypes.data_as(pd), vel.ctypes.data_as(pd), None if tp is None else tp.ctypes.data_as(pi), n,
                              None if lo is None else lo.ctypes.data_as(pll), 0 if lo is None else len(lo) - 1, period,
                              bsize, nsearch, nvel, approx, batch, den.ctypes.data_as(pd), stats.ctypes.data_as(pll))
        assert rc == 0, f"emulation failed with {rc}"
        return den, stats

    return run


def _agree(got, want):
    assert np.array_equal(got == 0, want == 0)
    nz = want != 0
    assert np.max(np.abs(got[nz] / want[nz] - 1.0)) < 1e-12
    return float(np.mean(got == want))


@pytest.mark.parametrize("approx", [1, 0])
def test_emulation_matches_oracle(emul, oracle, approx):
    rng = np.random.default_rng(5)
    n = 3000
    vel = rng.normal(size=(n, 3)) * 100
    clump = 50 + rng.normal(size=(n, 3)) * 0.5
    box = np.mod(50 + rng.normal(size=(n, 3)) * 30, 100.0)
    tied = (50 + rng.normal(size=(n, 3)) * 0.01).astype(np.float32).astype(np.float64)
    ptype = (rng.random(n) < 0.7).astype(np.int32)
    exact_bits = []
    for pos, kw in ((clump, {}), (box, dict(period=100.0)), (tied, {}), (box, dict(period=100.0, ptype=ptype)), (clump, dict(ptype=ptype)),
                    (clump, dict(nsearch=64, nvel=8, bsize=8)), (clump[:300], {})):
        want = oracle.velocity_density(pos, vel[:len(pos)], approx=approx, **kw)
        got, stats = emul(pos, vel[:len(pos)], approx=approx, **kw)
        assert stats[0] > len(pos) / kw.get("bsize", 16) - 1
        exact_bits.append(_agree(got, want))
    assert min(exact_bits) > 0.99


def test_periodic_wrapped_clump(emul, oracle):
    """a clump wrapped around the corner of the box: the reflected searches and their gates (Q16) decide the neighbour sets"""
    rng = np.random.default_rng(6)
    n = 4000
    pos = np.mod(rng.normal(size=(n, 3)) * 2.0, 100.0).astype(np.float32).astype(np.float64)
    vel = rng.normal(size=(n, 3)) * 50
    for approx in (1, 0):
        want = oracle.velocity_density(pos, vel, period=100.0, bsize=8, nsearch=64, nvel=8, approx=approx)
        got, _ = emul(pos, vel, period=100.0, bsize=8, nsearch=64, nvel=8, approx=approx)
        _agree(got, want)
        # and the choice matters: a minimum-image search gives other densities near the faces
        naive = oracle.velocity_density(np.mod(pos + 50.0, 100.0), vel, period=0.0, bsize=8, nsearch=64, nvel=8, approx=0) if approx == 0 else None
        if naive is not None:
            assert np.count_nonzero(np.abs(naive / want - 1.0) > 1e-6) > 0


def test_batches_and_golden(emul):
    """queries in several batches give the same densities; and the travelling fixture made from the compiled reference"""
    z = np.load(os.path.join(ROOT, "tests", "golden", "veldens", "vd_s9.npz"))
    pos, vel, ptype = z["pos"].astype(np.float64), z["vel"].astype(np.float64), z["ptype"].astype(np.int32)
    for approx, key in ((1, "den_approx"), (0, "den_exact")):
        whole, s1 = emul(pos, vel, approx=approx, ptype=ptype)
        parts, s2 = emul(pos, vel, approx=approx, ptype=ptype, batch=97)
        assert s1[1] == 1 and s2[1] > 5
        assert np.array_equal(whole, parts)
        _agree(whole, z[key])
    box = z["box"].astype(np.float64)
    for approx, key in ((1, "den_box_approx"), (0, "den_box_exact")):
        got, _ = emul(box, vel, period=100.0, bsize=8, nsearch=64, nvel=8, approx=approx, batch=500)
        _agree(got, z[key])


def caller_tree(oracle, pos, bsize):
    """what a caller that holds the reference's tree over the particles has: the tree order and the leaf boundaries"""
    perm, nodes = oracle.tree_order(pos, bsize)
    leaves = nodes[(nodes[:, 1] - nodes[:, 0]) <= bsize]
    leaves = leaves[np.argsort(leaves[:, 0])]
    assert leaves[0, 0] == 0 and np.array_equal(leaves[1:, 0], leaves[:-1, 1]) and leaves[-1, 1] == len(pos)
    return perm, np.concatenate([[0], leaves[:, 1]])


def test_leaves_of_the_callers_tree(emul, oracle):
    """SearchFullSet hands GetVelocityDensity its FOF tree (search.cxx:232): particles in tree order, leaves = that tree's.
    With tied coordinates a rebuilt tree would have other leaves; with the caller's leaf boundaries the densities are the
    reference's (which are the same whether it builds the tree itself or receives it: checked in test_oracle_vs_ref.py)."""
    rng = np.random.default_rng(21)
    n = 5000
    pos = (50 + rng.normal(size=(n, 3)) * 0.01).astype(np.float32).astype(np.float64)
    vel = rng.normal(size=(n, 3)) * 100
    ptype = (rng.random(n) < 0.8).astype(np.int32)
    want = oracle.velocity_density(pos, vel, ptype=ptype)
    perm, off = caller_tree(oracle, pos, 16)
    got, stats = emul(pos[perm], vel[perm], ptype=ptype[perm], leaf_offsets=off)
    assert stats[0] == len(off) - 1
    _agree(got, want[perm])
    # without the boundaries the rebuilt tree groups the tied particles differently: some densities move
    rebuilt, _ = emul(pos[perm], vel[perm], ptype=ptype[perm])
    assert np.count_nonzero(np.abs(rebuilt / np.where(want[perm] == 0, 1, want[perm]) - 1.0) > 1e-9) > 0
