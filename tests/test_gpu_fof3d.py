"""GPU parity: vrgpu_fof3d (through the C ABI) against the oracle restatement of KDTree::FOF.

Bar: bit-exact canonical membership (integer/index work).  The oracle itself is pinned against the compiled
reference in test_oracle_vs_ref.py / tests/golden.
"""
import os

import numpy as np
import pytest

import vrsynth

pytestmark = pytest.mark.gpu


def _b(nside, box=vrsynth.BOX):
    ell = box / nside
    # search.cxx:77: param[1] = ellxscale^2 * ellphys^2 * ellhalophysfac^2 ; fdist = sqrt(param[1])
    return np.sqrt((ell * ell) * (0.2 * 0.2) * (1.0 * 1.0))


@pytest.mark.parametrize("nside,seed", [(32, 7), (48, 11), (64, 7), (96, 5)])
def test_fof3d_periodic_matches_oracle(gpu_ctx, oracle, nside, seed):
    import vrgpu
    d = vrsynth.make_box(nside ** 3, seed)
    parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
    b = _b(nside)
    pf, ng = gpu_ctx.fof3d(parts, b, vrsynth.BOX, 20)
    po, no = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    assert ng == no
    # ids are canonical on both sides (size desc, min pid asc) so the arrays must be identical
    assert np.array_equal(pf, po)


def test_fof3d_open_box(gpu_ctx, oracle):
    import vrgpu
    d = vrsynth.make_box(40 ** 3, 3)
    parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
    b = _b(40)
    pf, ng = gpu_ctx.fof3d(parts, b, None, 10)
    po, no = oracle.fof3d(d["pos"], b, None, 10, d["pid"])
    assert ng == no and np.array_equal(pf, po)


def test_fof3d_known_answers(gpu_ctx):
    """Hand-made cases: strict '<' at exactly b, a chain across the periodic corner, empty / tiny inputs."""
    import vrgpu
    L, b = 100.0, 1.0
    # two particles at distance exactly b are NOT linked (strict <, KDLeafNode.cxx:558); at b - 1ulp they are
    pos = np.array([[10.0, 10.0, 10.0], [11.0, 10.0, 10.0], [0.0, 50.0, 50.0], [np.nextafter(1.0, 0.0), 50.0, 50.0]])
    parts = vrgpu.particles_from_arrays(pos)
    pf, ng = gpu_ctx.fof3d(parts, b, L, 2)
    assert ng == 1 and list(pf) == [0, 0, 1, 1]
    # chain through the corner of the periodic box
    pos = np.array([[99.7, 99.7, 99.7], [0.1, 0.1, 0.1], [0.6, 0.1, 0.1], [1.2, 0.1, 0.1], [40.0, 40.0, 40.0]])
    parts = vrgpu.particles_from_arrays(pos)
    pf, ng = gpu_ctx.fof3d(parts, b, L, 2)
    assert ng == 1 and list(pf) == [1, 1, 1, 1, 0]
    # same points, open box: the corner link disappears
    pf, ng = gpu_ctx.fof3d(parts, b, None, 2)
    assert ng == 1 and list(pf) == [0, 1, 1, 1, 0]
    # empty input
    parts = vrgpu.particles_from_arrays(np.zeros((0, 3)))
    pf, ng = gpu_ctx.fof3d(parts, b, L, 2)
    assert ng == 0 and pf.shape == (0,)


def test_fof3d_edge_cases(gpu_ctx, oracle):
    """Corners of the domain (tests/edge_cases.py): the oracle is pinned to the compiled reference on exactly these inputs
    (test_oracle_vs_ref.py::test_fof3d_edge_cases_vs_kdtree_fof)."""
    import edge_cases
    import vrgpu
    for name, (pos, period, minsize) in edge_cases.cases().items():
        pid = np.arange(len(pos))
        parts = vrgpu.particles_from_arrays(pos, pid=pid)
        pf, ng = gpu_ctx.fof3d(parts, edge_cases.B, period, minsize)
        want, no = oracle.fof3d(pos, edge_cases.B, period, minsize, pid)
        assert ng == no, name
        assert np.array_equal(pf, want), name


def test_fof3d_size_order_and_ties(gpu_ctx):
    """ids are 1..G by size descending with the min-pid tie break (BASELINE.json)."""
    import vrgpu
    rng = np.random.default_rng(1)
    clumps = []
    for c, k in ((10.0, 5), (30.0, 7), (50.0, 5), (70.0, 3)):
        clumps.append(c + 0.01 * rng.random((k, 3)))
    pos = np.concatenate(clumps)
    pid = np.arange(100, 100 + len(pos))[::-1].copy()  # reversed pids: tie between the two 5-clumps goes to the later one
    parts = vrgpu.particles_from_arrays(pos, pid=pid)
    pf, ng = gpu_ctx.fof3d(parts, 0.5, 100.0, 3)
    assert ng == 4
    assert set(pf[5:12]) == {1}
    assert set(pf[12:17]) == {2} and set(pf[0:5]) == {3}  # clump at 50 has the smaller pids
    assert set(pf[17:20]) == {4}


def test_fof3d_random_small_boxes(gpu_ctx, oracle):
    """120 random small inputs in boxes 5 to 24 linking lengths wide, periodic and open: most neighbour blocks wrap, the same
    block is reached through several periodic images, and the smallest boxes (no whole number of 4-cell blocks with
    2 c >= b) take the hash-table path of fof3d.cu instead of the block bitmap."""
    import vrgpu
    rng = np.random.default_rng(2027)
    for trial in range(120):
        b = float(rng.choice([0.5, 1.0, 1.7]))
        L = float(b * rng.uniform(4.7, 24.0))
        if int(np.ceil(L / (b / np.sqrt(3.0) * (1.0 - 1e-8)))) < 8:
            continue
        n = int(rng.integers(1, 600))
        centres = rng.random((int(rng.integers(1, 6)), 3)) * L
        pos = centres[rng.integers(0, len(centres), size=n)] + rng.normal(size=(n, 3)) * b * rng.uniform(0.2, 1.5)
        pos = np.mod(pos, L)
        if rng.random() < 0.2 and n > 3:
            pos[rng.integers(0, n, size=3), rng.integers(0, 3, size=3)] = rng.choice([0.0, L], size=3)
        pos = np.clip(pos.astype(np.float32).astype(np.float64), 0.0, L)
        periodic = rng.random() < 0.7
        minsize = int(rng.integers(1, 6))
        pid = np.arange(n)
        parts = vrgpu.particles_from_arrays(pos, pid=pid)
        pf, ng = gpu_ctx.fof3d(parts, b, L if periodic else None, minsize)
        want, no = oracle.fof3d(pos, b, L if periodic else None, minsize, pid)
        assert ng == no, (trial, n, b, L, periodic)
        assert np.array_equal(pf, want), (trial, n, b, L, periodic)


def test_fof3d_dense_cells_use_the_dense_queue(gpu_ctx, oracle):
    """Two well-filled cells two cells apart along x whose FIRST 4 x 4 particle pairs do not link while one later pair does:
    the short trial must leave the pair undecided and the warp-per-pair kernel must find the link."""
    import vrgpu
    rng = np.random.default_rng(3)
    L, b = 100.0, 1.0
    nc = int(np.ceil(np.ceil(L / (b / np.sqrt(3.0) * (1.0 - 1e-8))) / 4.0) * 4)
    c = L / nc
    yz = 17.5 * c + 0.01 * rng.random((82, 2))
    xa = np.concatenate([17.05 * c + 0.01 * rng.random(40), [17.9 * c]])
    xb = np.concatenate([19.8 * c + 0.01 * rng.random(40), [19.1 * c]])
    pos = np.column_stack([np.concatenate([xa, xb]), yz])
    pid = np.arange(len(pos))
    parts = vrgpu.particles_from_arrays(pos, pid=pid)
    pf, ng = gpu_ctx.fof3d(parts, b, L, 2)
    assert gpu_ctx.stats()["n_link_queue"] == 1
    want, no = oracle.fof3d(pos, b, L, 2, pid)
    assert ng == no == 1 and np.array_equal(pf, want) and np.all(pf == 1)
