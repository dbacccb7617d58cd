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


@pytest.mark.parametrize("form", ["keep_ids", "core_flags"])
def test_check_unbound_groups_as_search_sub_sub_calls_it(ref, refgpu, oracle, form):
    """The two other forms of CheckUnboundGroups on the hot path: ireorder = 0 (search.cxx:2867: ids
    kept, dissolved groups left with numingroup = 0) and ireorder = 1 with a group flag (search.cxx:2674: flagged groups first
    by size, the others by ascending size, flags permuted with the ids).  Direct sums only (every group below UNBINDNUM would be
    too small a sample, so the stock thresholds apply; groups above 150 take the tree on both sides)."""
    import vroracle
    nside = 40
    d = vrsynth.make_box(nside ** 3, 31)
    b = 0.2 * vrsynth.BOX / nside
    pf, ng = oracle.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
    pos = oracle.adjust_for_period(d["pos"], vrsynth.BOX, pf, ng)
    o = vroracle.default_opts(period=vrsynth.BOX, ellxscale=vrsynth.BOX / nside, fofbgtype=5, bound_halos=1, omp_fof=0)
    if form == "keep_ids":
        want = ref.unbind_full(pos, d["vel"], d["mass"], pf, ng, o, ireorder=0)
        got = refgpu.unbind_full(pos, d["vel"], d["mass"], pf, ng, o, ireorder=0)
    else:
        flag = np.ones(ng + 1, dtype=np.int64)
        flag[ng - ng // 3:] = 2                                   # the last third are "cores" (search.cxx:2671)
        want = ref.unbind_full(pos, d["vel"], d["mass"], pf, ng, o, ireorder=1, groupflag=flag)
        got = refgpu.unbind_full(pos, d["vel"], d["mass"], pf, ng, o, ireorder=1, groupflag=flag)
    assert got["changed"] == want["changed"] == 1
    # Each side against ITS oracle, exactly (as in test_unbinding_through_reference_types): the library against the intended
    # semantics of Q10, the reference against the oracle that carries oldnumingroup from group to group as the serial build does.
    # On this box the two differ by 40 bound flags and one surviving group (300 / 301).
    intended = oracle.unbind(pos, d["vel"], d["mass"], pf, ng, vroracle.unbind_opts_from(o, 150, sortflag_mode=0), d["pid"], 1e-6)
    carried = oracle.unbind(pos, d["vel"], d["mass"], pf, ng, vroracle.unbind_opts_from(o, 150, sortflag_mode=1, oldn_init=0), d["pid"], 1e-6)
    assert intended["n_marginal"] == 0
    for side, r, orc in (("library", got, intended), ("reference", want, carried)):
        bound = orc["pfof"] > 0
        assert r["ngroups"] == orc["ngroups"], side
        assert np.array_equal(r["pfof"] > 0, bound), side
        if form == "keep_ids":
            # ids kept, dissolved groups emptied, numingroup = the bound members left in every group
            assert np.array_equal(r["pfof"], np.where(bound, pf, 0)), side
            assert np.array_equal(r["numingroup"][1:], np.bincount(r["pfof"], minlength=ng + 1)[1:]), side
            assert np.count_nonzero(r["numingroup"][1:] > 0) == r["ngroups"], side
        else:
            gf, G = r["groupflag"], r["ngroups"]
            # the same partition of the bound particles as the oracle's (ids differ: another ordering rule)
            pairs = set(zip(r["pfof"][bound].tolist(), orc["pfof"][bound].tolist()))
            assert len(pairs) == G == len({a for a, _ in pairs}) == len({b for _, b in pairs}), side
            # flags travel with their groups
            assert np.array_equal(gf[r["pfof"][bound]], flag[pf[bound]]), side
            # flagged groups come first; inside each class the sizes are monotonic (descending, then ascending)
            num = np.bincount(r["pfof"], minlength=G + 1)   # (the reference leaves numingroup in the OLD numbering; its caller rebuilds it)
            cls = gf[1:G + 1]
            assert np.all(np.diff((cls != 1).astype(int)) >= 0), side
            n1 = num[1:G + 1][cls == 1]
            n2 = num[1:G + 1][cls != 1]
            assert np.all(np.diff(n1) <= 0) and np.all(np.diff(n2) >= 0), side
    nq10 = int(np.count_nonzero((got["pfof"] > 0) != (want["pfof"] > 0)))
    print(f"{form}: library {got['ngroups']} groups, reference {want['ngroups']}; Q10 particles {nq10}")
    assert nq10 == 40
