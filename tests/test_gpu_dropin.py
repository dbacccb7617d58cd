"""GPU: the drop-in boundary exercised through the reference's OWN types.  oracle/_ref/libvrref_gpu.so links the
reference's objects with velociraptor-stf_b200/host/vrgpu_adaptor.cxx (SearchFullSet / CheckUnboundGroups / Potential bodies on
top of the C ABI); the same bridge functions of libvrref.so run the reference's CPU bodies.  Same inputs, same Options."""
import os

import numpy as np
import pytest

import vrsynth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def refgpu():
    import vroracle
    if not vroracle.Ref.available("gpu") or not vroracle.Ref.available("serial"):
        pytest.skip("oracle/_ref harness not built")
    return vroracle.Ref("gpu")


@pytest.mark.parametrize("fofbgtype", [5, 3, 4])
def test_search_full_set_through_reference_types(ref, refgpu, fofbgtype):
    import vroracle
    nside = 40
    d = vrsynth.make_box(nside ** 3, 31)
    o = vroracle.default_opts(period=vrsynth.BOX, ellxscale=vrsynth.BOX / nside, fofbgtype=fofbgtype, omp_fof=0)
    want = ref.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], o)
    got = refgpu.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], o)
    cw, _ = vrsynth.canonical_groups(want["pfof"], d["pid"])
    assert got["ngroups"] == want["ngroups"]
    assert np.array_equal(got["pfof"], cw)          # the GPU ids are already canonical
    assert np.array_equal(got["pos"], want["pos"])  # periodic unwrapping + Q3 rounding written back into Particle[]
    assert np.array_equal(got["vel"], want["vel"])


def test_unbinding_through_reference_types(ref, refgpu, oracle):
    """Bound_halos=1 with the stock thresholds (direct sum n <= 150, kd-tree above) on a float32-exact box (tied kd-tree
    medians).  Two exact statements instead of a tolerance: (1) the library through the reference's types == the oracle with
    the INTENDED semantics of Q10 (oldnumingroup initialised per group), label for label; (2) the oracle with oldnumingroup
    carried from group to group, as the serial reference build happens to do, == the reference, label for label.  The
    particles on which (1) and (2) differ are Q10's (unbind.cxx:643,700 reads the variable uninitialised) and their number
    is pinned."""
    import vroracle
    nside = 40
    d = vrsynth.make_box(nside ** 3, 31)
    o = vroracle.default_opts(period=vrsynth.BOX, ellxscale=vrsynth.BOX / nside, fofbgtype=5, bound_halos=1, omp_fof=0)
    want = ref.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], o)
    got = refgpu.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], o)
    cw, _ = vrsynth.canonical_groups(want["pfof"], d["pid"])
    b = 0.2 * vrsynth.BOX / nside
    pf, ng = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    pos = oracle.adjust_for_period(d["pos"], vrsynth.BOX, pf, ng)
    intended = oracle.unbind(pos, d["vel"], d["mass"], pf, ng, vroracle.unbind_opts_from(o, 150, sortflag_mode=0), d["pid"], 1e-6)
    carried = oracle.unbind(pos, d["vel"], d["mass"], pf, ng, vroracle.unbind_opts_from(o, 150, sortflag_mode=1, oldn_init=0), d["pid"], 1e-6)
    ndiff = int(np.count_nonzero((got["pfof"] > 0) != (intended["pfof"] > 0)))
    assert ndiff <= intended["n_marginal"], f"{ndiff} bound flags differ from the oracle, {intended['n_marginal']} marginal"
    if ndiff == 0:
        assert got["ngroups"] == intended["ngroups"] and np.array_equal(got["pfof"], intended["pfof"])
    assert carried["ngroups"] == want["ngroups"] and np.array_equal(carried["pfof"], cw)
    nq10 = int(np.count_nonzero((intended["pfof"] > 0) != (cw > 0)))
    print(f"Q10 particles (intended vs carried oldnumingroup): {nq10} of {np.count_nonzero(cw > 0)} grouped")
    assert nq10 == 40


def test_potential_through_reference_types(ref, refgpu):
    import vroracle
    rng = np.random.default_rng(12)
    n = 20000
    r = rng.random(n) ** 1.5
    dd = rng.normal(size=(n, 3))
    pos = 50 + r[:, None] * dd / np.linalg.norm(dd, axis=1)[:, None]
    mass = 0.5 + rng.random(n)
    o = vroracle.default_opts(eps=0.01)
    want = ref.potential(pos, mass, o)
    got = refgpu.potential(pos, mass, o)
    assert np.max(np.abs(got / want - 1.0)) < 1e-10


def test_core_search_linking_through_reference_types(ref, refgpu):
    """Row f2, linking step: the adaptor body FOFCriterion6DCoreGPU receives the reference's own Particle array in the order
    its KDTree left it (ids = input index, potential field = the FOFcheckbg mask) and must return what
    tree->FOFCriterion(FOF6d, ..., FOFcheckbg) returns (search.cxx:1611), up to the canonical order of equal-sized groups."""
    n = 15000
    pos, vel, pot = vrsynth.halo_with_cores(n, 6)
    pid = np.arange(n)
    p6, p7 = vrsynth.core_search_lengths(pos, vel)
    for it in range(3):
        want, nw = ref.fofcriterion6d(pos, vel, pot, p6, p7, 0.5, 75, 16)
        got, ng = refgpu.fofcriterion6d(pos, vel, pot, p6, p7, 0.5, 75, 16)
        cw, _ = vrsynth.canonical_groups(want, pid)
        assert ng == nw and np.array_equal(got, cw)
        p6 *= 0.75 ** 2; p7 *= 0.75 ** 2
