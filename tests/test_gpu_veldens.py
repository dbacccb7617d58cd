"""GPU parity, SURVEY section 8 row f4: vrgpu_velocity_density (csrc/veldens.cu) against the oracle's restatement of
GetVelocityDensity (localfield.cxx:18), which is bit-identical to the compiled reference (tests/test_oracle_vs_ref.py), and
against the travelling fixture made from the reference.  Tolerance 1e-12 relative: every operation follows the reference's
order; the one exception is its pow(h, 3.0), cubed here with a single rounding.  The same per-thread source passes
these comparisons as a host emulation (tests/test_veldens_host.py); first green on the B200 in round 2."""
import os

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu]
HERE = os.path.dirname(os.path.abspath(__file__))


def _agree(got, want):
    assert np.array_equal(got == 0, want == 0)
    nz = want != 0
    assert np.max(np.abs(got[nz] / want[nz] - 1.0)) < 1e-12


@pytest.mark.parametrize("approx", [1, 0])
def test_velocity_density_matches_oracle(gpu_ctx, oracle, approx):
    import vrgpu
    rng = np.random.default_rng(5)
    n = 20000
    vel = rng.normal(size=(n, 3)) * 100
    clump = 50 + rng.normal(size=(n, 3)) * 0.5
    box = np.mod(50 + rng.normal(size=(n, 3)) * 30, 100.0)
    tied = (50 + rng.normal(size=(n, 3)) * 0.01).astype(np.float32).astype(np.float64)
    ptype = (rng.random(n) < 0.7).astype(np.int32)
    m = n if approx else 4000            # the oracle's per-particle search is quadratic
    for pos, kw in ((clump, {}), (box, dict(period=100.0)), (tied, {}), (box, dict(period=100.0, ptype=ptype)),
                    (clump, dict(nsearch=64, nvel=8, bsize=8)), (clump[:300], {})):
        pos = pos[:m]
        kw = {k: (v[:len(pos)] if k == "ptype" else v) for k, v in kw.items()}
        want = oracle.velocity_density(pos, vel[:len(pos)], approx=approx, **kw)
        parts = vrgpu.particles_from_arrays(pos, vel[:len(pos)])
        got = gpu_ctx.velocity_density(parts, approx=approx, **kw)
        _agree(got, want)


def test_velocity_density_golden(gpu_ctx):
    import vrgpu
    z = np.load(os.path.join(HERE, "golden", "veldens", "vd_s9.npz"))
    pos, vel, ptype = z["pos"].astype(np.float64), z["vel"].astype(np.float64), z["ptype"].astype(np.int32)
    parts = vrgpu.particles_from_arrays(pos, vel)
    for approx, key in ((1, "den_approx"), (0, "den_exact")):
        _agree(gpu_ctx.velocity_density(parts, approx=approx, ptype=ptype), z[key])
    box = vrgpu.particles_from_arrays(z["box"].astype(np.float64), vel)
    for approx, key in ((1, "den_box_approx"), (0, "den_box_exact")):
        _agree(gpu_ctx.velocity_density(box, period=100.0, bsize=8, nsearch=64, nvel=8, approx=approx), z[key])


def test_velocity_density_on_the_callers_leaves(gpu_ctx, oracle):
    """SearchFullSet's call (search.cxx:232): particles in the order of the caller's tree, that tree's leaf boundaries passed on"""
    import vrgpu
    from test_veldens_host import caller_tree
    rng = np.random.default_rng(21)
    n = 20000
    pos = (50 + rng.normal(size=(n, 3)) * 0.01).astype(np.float32).astype(np.float64)
    vel = rng.normal(size=(n, 3)) * 100
    ptype = (rng.random(n) < 0.8).astype(np.int32)
    want = oracle.velocity_density(pos, vel, ptype=ptype)
    perm, off = caller_tree(oracle, pos, 16)
    got = gpu_ctx.velocity_density(vrgpu.particles_from_arrays(pos[perm], vel[perm]), ptype=ptype[perm], leaf_offsets=off)
    _agree(got, want[perm])
    with pytest.raises(vrgpu.VrgpuError):
        gpu_ctx.velocity_density(vrgpu.particles_from_arrays(pos[perm], vel[perm]), leaf_offsets=off[:-1])   # does not reach n


def test_velocity_density_rejects_what_it_does_not_cover(gpu_ctx):
    import vrgpu
    rng = np.random.default_rng(1)
    parts = vrgpu.particles_from_arrays(rng.random((500, 3)), rng.random((500, 3)))
    with pytest.raises(vrgpu.VrgpuError):
        gpu_ctx.velocity_density(parts, nsearch=256, nvel=100)       # Nvel above the queue kept in registers
    with pytest.raises(vrgpu.VrgpuError):
        gpu_ctx.velocity_density(parts[:100], nsearch=256, nvel=32)  # fewer particles than Nsearch + 1


def test_velocity_density_through_reference_types(ref):
    """the adaptor body GetVelocityDensityGPU behind the reference's own Options / Particle types"""
    import vroracle
    if not vroracle.Ref.available("gpu"):
        pytest.skip("oracle/_ref harness not built")
    refgpu = vroracle.Ref("gpu")
    rng = np.random.default_rng(7)
    n = 15000
    pos = 50 + rng.normal(size=(n, 3)) * 0.5
    vel = rng.normal(size=(n, 3)) * 100
    ptype = (rng.random(n) < 0.8).astype(np.int32)
    for approx in (1, 0, 2):          # 2: the caller builds the tree and hands it in (search.cxx:232)
        want = ref.velocity_density(pos, vel, approx=approx, ptype=ptype)
        got = refgpu.velocity_density(pos, vel, approx=approx, ptype=ptype)
        _agree(got, want)
