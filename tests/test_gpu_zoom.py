"""GPU parity on the zoom-style input of BASELINE.json configs[4], scaled down: ONE 3D group holding (almost) every
particle, so the large-group branches are the ones exercised -- 6DFOF of a single big group with its own dispersion
(search.cxx:472-499, 520-579), kd-tree potentials (unbind.cxx:1607-1741) and the one-pass removal of groups above
ompunbindnum (unbind.cxx:552-639)."""
import os

import numpy as np
import pytest

import vrsynth

pytestmark = pytest.mark.gpu


def _inputs(n, seed=105):
    d = vrsynth.make_box(n, seed, single_halo=True)
    ell = vrsynth.BOX / float(n) ** (1.0 / 3.0)
    return d, ell


@pytest.mark.parametrize("fofbgtype,pp", [(3, 150), (5, 150), (3, -1)])
def test_single_halo_search_full_set(gpu_ctx, oracle, fofbgtype, pp):
    import vrgpu
    import vroracle
    n = 40000
    d, ell = _inputs(n)
    param1 = (ell * ell) * (0.2 * 0.2) * (1.0 * 1.0)
    b = float(np.sqrt(param1))
    pf, ng = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    assert ng == 1 and np.count_nonzero(pf == 1) > 0.95 * n          # the config's premise: one 3D group
    pos, vel = oracle.adjust_for_period(d["pos"], vrsynth.BOX, pf, ng), d["vel"]
    if fofbgtype == 3:
        o6 = oracle.fof6d(pos, vel, d["mass"], pf, ng, b, 1.25, 3, 20, d["pid"])
        pos, vel, pf, ng = o6["pos"], o6["vel"], o6["pfof"], o6["ngroups"]
        assert ng > 1                                                 # subclumps separate in phase space
    uo = vroracle.unbind_opts_from(vroracle.default_opts(), pp_threshold=pp, sortflag_mode=0)
    want = oracle.unbind(pos, vel, d["mass"], pf, ng, uo, d["pid"], 1e-6)
    parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
    so = vrgpu.search_opts(vrsynth.BOX, param1, halo_min_size=20, fofbgtype=fofbgtype, bound_halos=1,
                           unbind=vrgpu.default_unbind_opts(pp_threshold=pp))
    got, ngot = gpu_ctx.search_full_set(parts, so)
    st = gpu_ctx.stats()
    if pp > 0:
        assert st["n_tree_nodes"] > 0
    ndiff = int(np.count_nonzero((got > 0) != (want["pfof"] > 0)))
    assert ndiff <= want["n_marginal"] + st["n_marginal"]
    if ndiff == 0:
        assert ngot == want["ngroups"] and np.array_equal(got, want["pfof"])
    assert np.array_equal(parts["pos"], pos) and np.array_equal(parts["vel"], vel)




def test_single_halo_of_a_million_members_against_the_compiled_reference(gpu_ctx, ref, oracle):
    """BASELINE configs[4] at a size the CPU side still finishes in a minute: 1.2e6 particles in ONE 3D group (964 186 of them stay in
    the main 6D group), the stock settings.  3DFOF + 6DFOF: canonical membership and the rewritten positions / velocities bit for bit
    against the reference's own SearchFullSet -- the dispersion of the million-member group goes through the many-window evaluation
    (csrc/seqsum.h) and every bit of it enters the scaled coordinates.  Unbinding (kd-tree potentials of a 10^6-member tree, one-pass
    removal): against the oracle chained from the reference's 6DFOF output, bound sets identical up to the reported marginal
    particles (the reference itself adds quirk Q10 on the small subclumps, pinned in tests/test_gpu_dropin.py)."""
    import vrgpu
    import vroracle
    n = 1_200_000
    d, ell = _inputs(n)
    param1 = (ell * ell) * (0.2 * 0.2) * (1.0 * 1.0)
    o = vroracle.default_opts(period=vrsynth.BOX, ellxscale=ell, fofbgtype=3, bound_halos=0, omp_fof=0)
    want6 = ref.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], o)
    cw6, sizes6 = vrsynth.canonical_groups(want6["pfof"], d["pid"])
    assert sizes6[0] > 0.75 * n
    parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
    so = vrgpu.search_opts(vrsynth.BOX, param1, halo_min_size=20, fofbgtype=3, bound_halos=0)
    got6, ng6 = gpu_ctx.search_full_set(parts, so)
    assert ng6 == len(sizes6) == want6["ngroups"] and np.array_equal(got6, cw6)
    assert np.array_equal(parts["pos"], want6["pos"]) and np.array_equal(parts["vel"], want6["vel"])
    uo = vroracle.unbind_opts_from(vroracle.default_opts(period=vrsynth.BOX, ellxscale=ell, fofbgtype=3, bound_halos=1), 150, sortflag_mode=0)
    want = oracle.unbind(want6["pos"], want6["vel"], d["mass"], cw6, len(sizes6), uo, d["pid"], 1e-6)
    parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
    so = vrgpu.search_opts(vrsynth.BOX, param1, halo_min_size=20, fofbgtype=3, bound_halos=1, unbind=vrgpu.default_unbind_opts(pp_threshold=150))
    got, ng = gpu_ctx.search_full_set(parts, so)
    st = gpu_ctx.stats()
    ndiff = int(np.count_nonzero((got > 0) != (want["pfof"] > 0)))
    print(f"single halo of {n}: {ng6} 6D groups (largest {sizes6[0]}), {ng} bound groups (oracle {want['ngroups']}), bound flags differing {ndiff}, "
          f"marginal gpu {st['n_marginal']} / oracle {want['n_marginal']}, tree interactions {st['n_tree_interactions']}")
    assert ndiff <= want["n_marginal"] + st["n_marginal"]
    if ndiff == 0:
        assert ng == want["ngroups"] and np.array_equal(got, want["pfof"])


@pytest.mark.parametrize("pattern", ["zero_mean", "zero_axis", "ties", "equal_mass_drift", "mixed_mass"])
@pytest.mark.parametrize("multi_min", [0, 8192])
def test_big_group_dispersion_is_bit_exact(gpu_ctx, oracle, pattern, multi_min, monkeypatch):
    """multi_min = 8192 sends the group through the many-window evaluation (VRGPU_DISP_MULTI_MIN, csrc/seqsum.h "Many windows
    at once": summaries under a predicted unit + one walker) -- the same bit-exact bar.

    The velocity dispersion of a large group is a left-to-right fp64 sum in the reference (search.cxx:480-493); groups
    above 4096 members take the integer-scan evaluation (csrc/seqsum.h).  Membership and the scaled/unscaled coordinates
    (Q3) depend on every bit of it, so comparing those with the oracle pins the sum."""
    import vrgpu
    if multi_min:
        monkeypatch.setenv("VRGPU_DISP_MULTI_MIN", str(multi_min))
    else:
        monkeypatch.delenv("VRGPU_DISP_MULTI_MIN", raising=False)
    n = 30000
    rng = np.random.default_rng({"zero_mean": 1, "zero_axis": 2, "ties": 3, "equal_mass_drift": 4, "mixed_mass": 5}[pattern])
    pos = rng.normal(size=(n, 3)) * 0.3 + 50.0
    vel = rng.normal(size=(n, 3)) * 200.0
    mass = np.full(n, float(np.float32(0.0207)))
    if pattern == "zero_axis":
        vel[:, 0] = 0.0
    elif pattern == "ties":
        vel = np.round(vel * 2.0) / 2.0
        mass[:] = 0.5
    elif pattern == "equal_mass_drift":
        vel += np.array([700.0, -300.0, 50.0])
    elif pattern == "mixed_mass":
        mass = mass * (1.0 + (rng.random(n) < 0.1) * 7.0) * rng.choice([1.0, 1e-3, 1e3], size=n)
    vel = vel.astype(np.float32).astype(np.float64)
    pf3 = np.ones(n, dtype=np.int64)
    pid = np.arange(1, n + 1, dtype=np.int64)
    for mode in (3, 4):
        want = oracle.fof6d(pos, vel, mass, pf3, 1, 0.05, 1.25, mode, 20, pid)
        parts = vrgpu.particles_from_arrays(pos, vel, mass, pid)
        got, ng = gpu_ctx.fof6d(parts, pf3, 1, 0.05, 1.25, mode, 20)
        assert ng == want["ngroups"] and np.array_equal(got, want["pfof"])
        assert np.array_equal(parts["vel"], want["vel"]) and np.array_equal(parts["pos"], want["pos"])
