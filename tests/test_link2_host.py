"""CPU: the candidate-queue variant of the 3D cell linker (velociraptor-stf_b200/csrc/link2.cuh, VRGPU_LINK_VARIANT=1)
run as a serial emulation -- the SAME per-thread source the CUDA kernels call, compiled by g++ -- against the oracle's
KDTree::FOF restatement.  Bar: identical canonical membership, for every emulated thread order."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import vrsynth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("link2") / "liblink2_emul.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-Wall", "-I",
                    os.path.join(ROOT, "velociraptor-stf_b200", "csrc"), os.path.join(ROOT, "tests", "host", "link2_emul.cpp"),
                    "-o", so], check=True)
    lib = C.CDLL(so)
    lib.link2_emul.restype = C.c_int
    lib.link2_emul.argtypes = [C.POINTER(C.c_double), C.c_longlong, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int,
                               C.c_uint, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]

    def run(pos, b, period, order_mode=0, seed=1, axes=(1, 1, 1), filter_pass2=1):
        pos = np.ascontiguousarray(pos, dtype=np.float64)
        n = pos.shape[0]
        root = np.zeros(n, dtype=np.int64)
        stats = np.zeros(7, dtype=np.int64)
        rc = lib.link2_emul(pos.ctypes.data_as(C.POINTER(C.c_double)), n, b, period or 0.0, *axes, order_mode, seed, filter_pass2,
                            root.ctypes.data_as(C.POINTER(C.c_longlong)), stats.ctypes.data_as(C.POINTER(C.c_longlong)))
        assert rc == 0, f"emulation failed with {rc}"
        return root, stats

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


@pytest.mark.parametrize("nside,seed", [(32, 7), (48, 11), (64, 7)])
@pytest.mark.parametrize("order_mode", [0, 1, 2])
def test_link2_periodic_matches_oracle(emul, oracle, nside, seed, order_mode):
    d = vrsynth.make_box(nside ** 3, seed)
    b = _b(nside)
    root, stats = emul(d["pos"], b, vrsynth.BOX, order_mode, seed)
    got, ng = canonical(root, d["pid"], 20)
    want, no = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    assert ng == no
    assert np.array_equal(got, want)
    assert stats[1] > 0 and stats[5] > 0


def test_link2_unfiltered_second_pass(emul, oracle):
    """VRGPU_LINK_VARIANT=2: pass 2 emits every second-shell neighbour and leaves the root check to the pair threads."""
    d = vrsynth.make_box(48 ** 3, 5)
    b = _b(48)
    root, stats = emul(d["pos"], b, vrsynth.BOX, 2, 5, filter_pass2=0)
    rootf, statsf = emul(d["pos"], b, vrsynth.BOX, 2, 5, filter_pass2=1)
    assert stats[2] > 10 * statsf[2]                      # the filter removes most of the second-shell candidates
    got, ng = canonical(root, d["pid"], 20)
    want, no = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    assert ng == no and np.array_equal(got, want)
    assert np.array_equal(canonical(rootf, d["pid"], 20)[0], want)


def test_link2_open_box(emul, oracle):
    d = vrsynth.make_box(40 ** 3, 3)
    b = _b(40)
    root, _ = emul(d["pos"], b, None, 2, 5)
    got, ng = canonical(root, d["pid"], 10)
    want, no = oracle.fof3d(d["pos"], b, None, 10, d["pid"])
    assert ng == no and np.array_equal(got, want)


def test_link2_dense_cells_use_the_dense_queue(emul, oracle):
    """Two well-filled cells two cells apart along x whose FIRST 4 x 4 particle pairs do not link while one later pair
    does: the inline trial must leave the pair undecided and the exhaustive pass must find the link."""
    rng = np.random.default_rng(3)
    L, b = 100.0, 1.0
    nc = int(np.ceil(L / (b / np.sqrt(3.0) * (1.0 - 1e-8))))
    c = L / nc                                            # 0.5747: cell 17 = [9.770, 10.345), cell 19 = [10.92, 11.49)
    yz = 17.5 * c + 0.01 * rng.random((82, 2))
    xa = np.concatenate([17.05 * c + 0.01 * rng.random(40), [17.9 * c]])   # bulk of A at 9.80, its last particle at 10.29
    xb = np.concatenate([19.8 * c + 0.01 * rng.random(40), [19.1 * c]])    # bulk of B at 11.38, its last particle at 10.98
    pos = np.column_stack([np.concatenate([xa, xb]), yz])
    pid = np.arange(len(pos))
    for mode in (0, 1, 2):
        root, stats = emul(pos, b, L, mode, 9)
        assert stats[4] == 1 and stats[6] == 1            # one dense pair in pass 2, united by the exhaustive test
        got, ng = canonical(root, pid, 2)
        want, no = oracle.fof3d(pos, b, L, 2, pid)
        assert ng == no == 1 and np.array_equal(got, want) and np.all(got == 1)


def test_link2_edge_cases(emul, oracle):
    """Corners of the domain (tests/edge_cases.py; the oracle is pinned to the compiled reference on the same inputs in
    test_oracle_vs_ref.py): coincident particles, particles on the faces, minsize boundaries, one cell, corner chains."""
    import edge_cases
    for name, (pos, period, minsize) in edge_cases.cases().items():
        pid = np.arange(len(pos))
        want, no = oracle.fof3d(pos, edge_cases.B, period, minsize, pid)
        for mode in (0, 1, 2):
            root, _ = emul(pos, edge_cases.B, period, mode, 3)
            got, ng = canonical(root, pid, minsize)
            assert ng == no, name
            assert np.array_equal(got, want), name


def test_link2_known_answers(emul):
    L, b = 100.0, 1.0
    pos = np.array([[10.0, 10.0, 10.0], [11.0, 10.0, 10.0], [0.0, 50.0, 50.0], [np.nextafter(1.0, 0.0), 50.0, 50.0]])
    root, _ = emul(pos, b, L)
    pf, ng = canonical(root, np.arange(4), 2)
    assert ng == 1 and list(pf) == [0, 0, 1, 1]
    pos = np.array([[99.7, 99.7, 99.7], [0.1, 0.1, 0.1], [0.6, 0.1, 0.1], [1.2, 0.1, 0.1], [40.0, 40.0, 40.0]])
    root, _ = emul(pos, b, L)
    pf, ng = canonical(root, np.arange(5), 2)
    assert ng == 1 and list(pf) == [1, 1, 1, 1, 0]
    root, _ = emul(pos, b, None)
    pf, ng = canonical(root, np.arange(5), 2)
    assert ng == 1 and list(pf) == [0, 1, 1, 1, 0]


def test_link2_slab_open_along_x(emul, oracle):
    """A multi-GPU slab links with x open and y, z periodic (multigpu.cu): compare with the oracle on a box whose
    particles keep clear of the x faces, so that the open x axis cannot cut a group."""
    d = vrsynth.make_box(32 ** 3, 13)
    b = _b(32)
    pos = d["pos"].copy()
    keep = (pos[:, 0] > 2 * b) & (pos[:, 0] < vrsynth.BOX - 2 * b)
    pos = pos[keep]
    pid = np.arange(len(pos))
    root, _ = emul(pos, b, vrsynth.BOX, 2, 3, axes=(0, 1, 1))
    got, ng = canonical(root, pid, 20)
    want, no = oracle.fof3d(pos, b, vrsynth.BOX, 20, pid)
    assert ng == no and np.array_equal(got, want)


def test_link2_random_small_inputs(emul, oracle):
    """300 random small inputs: clustered points in boxes only a few linking lengths wide (so that most neighbour rows wrap
    and the same cell is reached through several periodic images), periodic and open, every emulated thread order."""
    rng = np.random.default_rng(2026)
    for trial in range(300):
        b = float(rng.choice([0.5, 1.0, 1.7]))
        L = float(b * rng.uniform(4.7, 14.0))          # 8.1 .. 24 fine cells per axis (the library needs at least 8)
        nc = int(np.ceil(L / (b / np.sqrt(3.0) * (1.0 - 1e-8))))
        if nc < 8:
            continue
        n = int(rng.integers(1, 400))
        centres = rng.random((int(rng.integers(1, 6)), 3)) * L
        pos = centres[rng.integers(0, len(centres), size=n)] + rng.normal(size=(n, 3)) * b * rng.uniform(0.2, 1.5)
        if rng.random() < 0.3:                          # some exactly coincident particles and some on the faces
            pos[: n // 4] = pos[n // 4: 2 * (n // 4)][: n // 4] if n >= 8 else pos[: n // 4]
        pos = np.mod(pos, L)
        if rng.random() < 0.2 and n > 3:
            pos[rng.integers(0, n, size=3), rng.integers(0, 3, size=3)] = rng.choice([0.0, L], size=3)
        pos = pos.astype(np.float32).astype(np.float64)
        pos = np.clip(pos, 0.0, L)
        periodic = rng.random() < 0.7
        minsize = int(rng.integers(1, 6))
        pid = np.arange(n)
        want, no = oracle.fof3d(pos, b, L if periodic else None, minsize, pid)
        root, _ = emul(pos, b, L if periodic else None, int(trial % 3), trial)
        got, ng = canonical(root, pid, minsize)
        assert ng == no, (trial, n, b, L, periodic)
        assert np.array_equal(got, want), (trial, n, b, L, periodic)
