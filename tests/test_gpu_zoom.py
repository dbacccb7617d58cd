"""GPU parity on the zoom-style input of BASELINE.json configs[4], scaled down: ONE 3D group holding (almost) every
particle, so the large-group branches are the ones exercised -- 6DFOF of a single big group with its own dispersion
(search.cxx:472-499, 520-579), kd-tree potentials (unbind.cxx:1607-1741) and the one-pass removal of groups above
ompunbindnum (unbind.cxx:552-639)."""
import numpy as np
import pytest

import vrsynth

pytestmark = pytest.mark.gpu


def _inputs(n, seed=105):
    d = vrsynth.make_box(n, seed, single_halo=True)
    # de-quantise: float32-exact coordinates tie at kd-tree medians (see test_gpu_tree.test_tree_with_tied_coordinates)
    rng = np.random.default_rng(2)
    d["pos"] = np.mod(d["pos"] + rng.random(d["pos"].shape) * 1e-7, vrsynth.BOX)
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
