"""GPU parity: 6DFOF inside each 3D group (vrgpu_fof6d / vrgpu_search_full_set with FoF_Field_search_type 3 and 4)
against the oracle restatement of search.cxx:400-698 and against the reference's golden vectors.

Bar: bit-exact canonical membership AND bit-exact phase-space coordinates after the scale/unscale round trip (Q3)."""
import glob
import os

import numpy as np
import pytest

import vrsynth

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))


def _b(nside):
    ell = vrsynth.BOX / nside
    return float(np.sqrt((ell * ell) * (0.2 * 0.2) * (1.0 * 1.0)))


@pytest.mark.parametrize("nside,seed,mode", [(32, 11, 3), (48, 5, 3), (64, 9, 3), (32, 11, 4), (48, 5, 4)])
def test_fof6d_matches_oracle(gpu_ctx, oracle, nside, seed, mode):
    import vrgpu
    d = vrsynth.make_box(nside ** 3, seed)
    b = _b(nside)
    pf3, ng3 = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    pos = oracle.adjust_for_period(d["pos"], vrsynth.BOX, pf3, ng3)
    want = oracle.fof6d(pos, d["vel"], d["mass"], pf3, ng3, b, 1.25, mode, 20, d["pid"])
    parts = vrgpu.particles_from_arrays(pos, d["vel"], d["mass"], d["pid"])
    got, ng = gpu_ctx.fof6d(parts, pf3, ng3, b, 1.25, mode, 20)
    assert ng == want["ngroups"]
    assert np.array_equal(got, want["pfof"])
    assert np.array_equal(parts["pos"], want["pos"]) and np.array_equal(parts["vel"], want["vel"])


def test_fof6d_small_cases(gpu_ctx, oracle):
    """Two clumps at the same place separated only in velocity; a clump below the size cut; unequal masses."""
    import vrgpu
    rng = np.random.default_rng(4)
    n1, n2, n3 = 400, 300, 12
    pos = 50 + 0.05 * rng.normal(size=(n1 + n2 + n3, 3))
    vel = np.concatenate([rng.normal(0, 10, (n1, 3)), rng.normal(0, 10, (n2, 3)) + [400.0, 0, 0], rng.normal(0, 5, (n3, 3)) - [0, 900.0, 0]])
    mass = 0.5 + rng.random(len(pos))
    pf3 = np.ones(len(pos), dtype=np.int64)
    want = oracle.fof6d(pos, vel, mass, pf3, 1, 0.2, 0.3, 3, 20)
    parts = vrgpu.particles_from_arrays(pos, vel, mass)
    got, ng = gpu_ctx.fof6d(parts, pf3, 1, 0.2, 0.3, 3, 20)
    assert want["ngroups"] >= 2
    assert ng == want["ngroups"] and np.array_equal(got, want["pfof"])
    # no 3D groups at all
    got, ng = gpu_ctx.fof6d(parts, np.zeros(len(pos), dtype=np.int64), 0, 0.2, 0.3, 3, 20)
    assert ng == 0 and not got.any()


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_search_full_set_6dfof_golden(gpu_ctx, path):
    """SearchFullSet with FoF_Field_search_type=3 in one resident call vs the reference's own output."""
    import vrgpu
    z = np.load(path)
    n = z["pos"].shape[0]
    nside = int(z["nside"])
    parts = vrgpu.particles_from_arrays(z["pos"].astype(np.float64), z["vel"].astype(np.float64), np.full(n, np.float64(z["mass"])),
                                        np.arange(1, n + 1, dtype=np.int64))
    ell = float(z["box"]) / nside
    so = vrgpu.search_opts(float(z["box"]), (ell * ell) * (0.2 * 0.2) * (1.0 * 1.0), halo_min_size=20, fofbgtype=3)
    got, ng = gpu_ctx.search_full_set(parts, so)
    assert ng == int(z["ng6d"])
    assert np.array_equal(got, z["pfof6d"].astype(np.int64))
    so5 = vrgpu.search_opts(float(z["box"]), (ell * ell) * (0.2 * 0.2) * (1.0 * 1.0), halo_min_size=20, fofbgtype=5)
    parts = vrgpu.particles_from_arrays(z["pos"].astype(np.float64), z["vel"].astype(np.float64), np.full(n, np.float64(z["mass"])),
                                        np.arange(1, n + 1, dtype=np.int64))
    got, ng = gpu_ctx.search_full_set(parts, so5)
    assert ng == int(z["ng3d"]) and np.array_equal(got, z["pfof3d"].astype(np.int64))
    assert np.array_equal(parts["pos"], z["pos_out"])


def test_search_full_set_6dfof_unbind(gpu_ctx, oracle):
    """type 3 + Bound_halos=1: 3DFOF -> 6DFOF -> CheckUnboundGroups, resident between the stages."""
    import vrgpu
    import vroracle
    nside = 48
    d = vrsynth.make_box(nside ** 3, 13)
    b = _b(nside)
    pf3, ng3 = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    pos = oracle.adjust_for_period(d["pos"], vrsynth.BOX, pf3, ng3)
    o6 = oracle.fof6d(pos, d["vel"], d["mass"], pf3, ng3, b, 1.25, 3, 20, d["pid"])
    uo = vroracle.unbind_opts_from(vroracle.default_opts(), pp_threshold=-1, sortflag_mode=0)
    want = oracle.unbind(o6["pos"], o6["vel"], d["mass"], o6["pfof"], o6["ngroups"], uo, d["pid"], 1e-6)
    parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
    ell = vrsynth.BOX / nside
    so = vrgpu.search_opts(vrsynth.BOX, (ell * ell) * (0.2 * 0.2) * (1.0 * 1.0), halo_min_size=20, fofbgtype=3, bound_halos=1,
                           unbind=vrgpu.default_unbind_opts(pp_threshold=-1))
    got, ng = gpu_ctx.search_full_set(parts, so)
    ndiff = int(np.count_nonzero((got > 0) != (want["pfof"] > 0)))
    assert ndiff <= want["n_marginal"] + gpu_ctx.stats()["n_marginal"]
    if ndiff == 0:
        assert ng == want["ngroups"] and np.array_equal(got, want["pfof"])
    assert np.array_equal(parts["pos"], o6["pos"]) and np.array_equal(parts["vel"], o6["vel"])


def test_pinned_caller_memory_gets_identical_writeback(gpu_ctx):
    """Page-locked caller records take the zero-copy write-back kernels; pageable ones the staged host scatter."""
    import vrgpu
    nside = 40
    d = vrsynth.make_box(nside ** 3, 17)
    ell = vrsynth.BOX / nside
    so = vrgpu.search_opts(vrsynth.BOX, (ell * ell) * (0.2 * 0.2) * (1.0 * 1.0), halo_min_size=20, fofbgtype=3, bound_halos=1)
    a = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"], pinned=False)
    b = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"], pinned=True)
    pa, na = gpu_ctx.search_full_set(a, so)
    pb, nb = gpu_ctx.search_full_set(b, so)
    assert na == nb and np.array_equal(pa, pb)
    assert np.array_equal(a["pos"], b["pos"]) and np.array_equal(a["vel"], b["vel"])
    assert not np.array_equal(a["pos"], d["pos"])  # something was unwrapped / re-rounded


def test_core_search_linking(gpu_ctx, oracle):
    """SURVEY section 8 row f2, linking step: vrgpu_fof_criterion6d against the travelling fixture made from the reference's
    KDTree::FOFCriterion(FOF6d, ..., FOFcheckbg) (tests/golden/core) and against the oracle on two more halos, three
    iterations of shrinking linking lengths each.  Bar: identical canonical membership."""
    import vrgpu
    import vrsynth
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "core", "core_s3.npz"))
    pos, vel, pot = z["pos"].astype(np.float64), z["vel"].astype(np.float64), z["pot"]
    parts = vrgpu.particles_from_arrays(pos, vel, pid=np.arange(len(pos)))
    for it in range(3):
        pf, ng = gpu_ctx.fof_criterion6d(parts, pot, float(z[f"p6_{it}"]), float(z[f"p7_{it}"]), float(z["p9"]), int(z["minsize"]))
        assert ng == int(z[f"ng_{it}"])
        assert np.array_equal(pf, z[f"pfof_{it}"].astype(np.int64))
    for n, seed in ((6000, 1), (40000, 4)):
        pos, vel, pot = vrsynth.halo_with_cores(n, seed)
        pid = np.arange(n)
        parts = vrgpu.particles_from_arrays(pos, vel, pid=pid)
        p6, p7 = vrsynth.core_search_lengths(pos, vel)
        for it in range(3):
            minsize = max(20, int(n * 0.005))
            want, nw = oracle.fof_criterion6d(pos, vel, pot, p6, p7, 0.5, minsize, pid)
            got, ng = gpu_ctx.fof_criterion6d(parts, pot, p6, p7, 0.5, minsize)
            assert ng == nw and np.array_equal(got, want)
            p6 *= 0.75 ** 2; p7 *= 0.75 ** 2
    # no mask, and nobody below the threshold
    got, ng = gpu_ctx.fof_criterion6d(parts, None, p6, p7, 0.5, 20)
    want, nw = oracle.fof_criterion6d(pos, vel, np.zeros(n), p6, p7, 0.5, 20, pid)
    assert ng == nw and np.array_equal(got, want)
    got, ng = gpu_ctx.fof_criterion6d(parts, np.ones(n), p6, p7, 0.5, 20)
    assert ng == 0 and not got.any()
