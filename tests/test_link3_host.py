"""CPU: the neighbour enumeration of the 3D cell linker (velociraptor-stf_b200/csrc/link3.cuh: block-major cell keys, block
occupancy masks, bitmap + rank look-up, (local index, block) candidate tables) compiled by g++ -- the SAME source the CUDA
kernels of fof3d.cu call -- (1) against a brute-force walk over the 62 half-space offsets and (2) as a serial linker against
the oracle's KDTree::FOF restatement.  Bar: identical canonical membership."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import vrsynth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("link3") / "liblink3_emul.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-Wall", "-I",
                    os.path.join(ROOT, "velociraptor-stf_b200", "csrc"), os.path.join(ROOT, "tests", "host", "link3_emul.cpp"),
                    "-o", so], check=True)
    L = C.CDLL(so)
    L.link3_check_tables.restype = C.c_longlong
    L.link3_fof.restype = C.c_longlong
    L.link3_fof.argtypes = [C.POINTER(C.c_double), C.c_longlong, C.c_double, C.c_double, C.POINTER(C.c_int), C.c_int,
                            C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
    return L


@pytest.fixture(scope="module")
def emul(lib):
    def run(pos, b, period, order_mode=0, axes=(1, 1, 1)):
        pos = np.ascontiguousarray(pos, dtype=np.float64)
        n = pos.shape[0]
        lab = np.zeros(n, dtype=np.int64)
        stats = np.zeros(4, dtype=np.int64)
        ax = (C.c_int * 3)(*axes)
        rc = lib.link3_fof(pos.ctypes.data_as(C.POINTER(C.c_double)), n, b, period or 0.0, ax, order_mode,
                           lab.ctypes.data_as(C.POINTER(C.c_longlong)), stats.ctypes.data_as(C.POINTER(C.c_longlong)))
        if rc == -5:
            return None, None     # box too small for whole blocks: the library's hash-table path, not this code
        assert rc >= 0, f"emulation failed with {rc}"
        return lab, stats
    return run


def canonical(root, pid, minsize):
    """ids 1..G by (size desc, min pid asc), 0 below minsize -- the library's rule (BASELINE.json)."""
    n = len(root)
    if n == 0:
        return np.zeros(0, dtype=np.int64), 0
    u, inv, cnt = np.unique(root, return_inverse=True, return_counts=True)
    minpid = np.full(len(u), np.iinfo(np.int64).max)
    np.minimum.at(minpid, inv, pid)
    keep = np.flatnonzero(cnt >= minsize)
    rank = keep[np.lexsort((minpid[keep], -cnt[keep]))]
    gid = np.zeros(len(u), dtype=np.int64)
    gid[rank] = np.arange(1, len(rank) + 1)
    return gid[inv], len(rank)


def _b(nside):
    return 0.2 * vrsynth.BOX / nside


def test_candidate_tables_are_the_62_half_space_offsets(lib):
    assert lib.link3_check_tables() == 0


@pytest.mark.parametrize("nside,seed", [(32, 7), (48, 11), (64, 7)])
@pytest.mark.parametrize("order_mode", [0, 1])
def test_link3_periodic_matches_oracle(emul, oracle, nside, seed, order_mode):
    d = vrsynth.make_box(nside ** 3, seed)
    b = _b(nside)
    lab, stats = emul(d["pos"], b, vrsynth.BOX, order_mode)
    got, ng = canonical(lab, d["pid"], 20)
    want, no = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    assert ng == no
    assert np.array_equal(got, want)
    assert stats[2] > 0 and stats[1] < stats[0]


def test_link3_open_box(emul, oracle):
    d = vrsynth.make_box(40 ** 3, 3)
    b = _b(40)
    lab, _ = emul(d["pos"], b, None, 1)
    got, ng = canonical(lab, d["pid"], 10)
    want, no = oracle.fof3d(d["pos"], b, None, 10, d["pid"])
    assert ng == no and np.array_equal(got, want)


def test_link3_edge_cases(emul, oracle):
    """Corners of the domain (tests/edge_cases.py; the oracle is pinned to the compiled reference on the same inputs in
    test_oracle_vs_ref.py): coincident particles, particles on the faces, minsize boundaries, one cell, corner chains."""
    import edge_cases
    for name, (pos, period, minsize) in edge_cases.cases().items():
        pid = np.arange(len(pos))
        want, no = oracle.fof3d(pos, edge_cases.B, period, minsize, pid)
        for mode in (0, 1):
            lab, _ = emul(pos, edge_cases.B, period, mode)
            got, ng = canonical(lab, pid, minsize)
            assert ng == no, name
            assert np.array_equal(got, want), name


def test_link3_known_answers(emul):
    L, b = 100.0, 1.0
    pos = np.array([[10.0, 10.0, 10.0], [11.0, 10.0, 10.0], [0.0, 50.0, 50.0], [np.nextafter(1.0, 0.0), 50.0, 50.0]])
    lab, _ = emul(pos, b, L)
    pf, ng = canonical(lab, np.arange(4), 2)
    assert ng == 1 and list(pf) == [0, 0, 1, 1]
    pos = np.array([[99.7, 99.7, 99.7], [0.1, 0.1, 0.1], [0.6, 0.1, 0.1], [1.2, 0.1, 0.1], [40.0, 40.0, 40.0]])
    lab, _ = emul(pos, b, L)
    pf, ng = canonical(lab, np.arange(5), 2)
    assert ng == 1 and list(pf) == [1, 1, 1, 1, 0]
    lab, _ = emul(pos, b, None)
    pf, ng = canonical(lab, np.arange(5), 2)
    assert ng == 1 and list(pf) == [0, 1, 1, 1, 0]


def test_link3_slab_open_along_x(emul, oracle):
    """A multi-GPU slab links with x open and y, z periodic (multigpu.cu): compare with the oracle on a box whose
    particles keep clear of the x faces, so that the open x axis cannot cut a group."""
    d = vrsynth.make_box(32 ** 3, 13)
    b = _b(32)
    pos = d["pos"].copy()
    keep = (pos[:, 0] > 2 * b) & (pos[:, 0] < vrsynth.BOX - 2 * b)
    pos = pos[keep]
    pid = np.arange(len(pos))
    lab, _ = emul(pos, b, vrsynth.BOX, 1, axes=(0, 1, 1))
    got, ng = canonical(lab, pid, 20)
    want, no = oracle.fof3d(pos, b, vrsynth.BOX, 20, pid)
    assert ng == no and np.array_equal(got, want)


def test_link3_random_small_inputs(emul, oracle):
    """300 random small inputs: clustered points in boxes only a few linking lengths wide (so that most neighbour blocks wrap
    and the same block is reached through several periodic images), periodic and open."""
    rng = np.random.default_rng(2026)
    ran = 0
    for trial in range(400):
        b = float(rng.choice([0.5, 1.0, 1.7]))
        L = float(b * rng.uniform(4.7, 24.0))          # 8.1 .. 42 fine cells per axis (the library needs at least 8)
        nc = int(np.ceil(L / (b / np.sqrt(3.0) * (1.0 - 1e-8))))
        if nc < 8:
            continue
        n = int(rng.integers(1, 400))
        centres = rng.random((int(rng.integers(1, 6)), 3)) * L
        pos = centres[rng.integers(0, len(centres), size=n)] + rng.normal(size=(n, 3)) * b * rng.uniform(0.2, 1.5)
        if rng.random() < 0.3:                          # some exactly coincident particles
            pos[: n // 4] = pos[n // 4: 2 * (n // 4)][: n // 4] if n >= 8 else pos[: n // 4]
        pos = np.mod(pos, L)
        if rng.random() < 0.2 and n > 3:                # and some on the faces
            pos[rng.integers(0, n, size=3), rng.integers(0, 3, size=3)] = rng.choice([0.0, L], size=3)
        pos = pos.astype(np.float32).astype(np.float64)
        pos = np.clip(pos, 0.0, L)
        periodic = rng.random() < 0.7
        minsize = int(rng.integers(1, 6))
        pid = np.arange(n)
        want, no = oracle.fof3d(pos, b, L if periodic else None, minsize, pid)
        lab, _ = emul(pos, b, L if periodic else None, int(trial % 2))
        if lab is None:
            continue
        ran += 1
        got, ng = canonical(lab, pid, minsize)
        assert ng == no, (trial, n, b, L, periodic)
        assert np.array_equal(got, want), (trial, n, b, L, periodic)
        lab1, _ = emul(pos, b, L if periodic else None, 0)
        assert np.array_equal(canonical(lab1, pid, 1)[0], oracle.fof3d(pos, b, L if periodic else None, 1, pid)[0]), (trial, "minsize 1")
    assert ran > 250
