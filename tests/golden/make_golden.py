"""Generate tests/golden/*.npz from the UNMODIFIED reference compiled into oracle/_ref (run where
/root/reference exists: `python tests/golden/make_golden.py`).  The vectors pin the oracle restatement and
the CUDA path on machines where the reference cannot be rebuilt.

Each file holds the float32-exact inputs and the reference's outputs:
  pfof3d     SearchFullSet, FoF_Field_search_type=5 (3DFOF), Bound_halos=0, canonical ids
  pfof6d     SearchFullSet, FoF_Field_search_type=3 (adaptive 6DFOF inside each 3D group), canonical ids
  pfof_unb   SearchFullSet, type 5 + Bound_halos=1 (CheckUnboundGroups: direct sum n<=150, tree above)
  phi_tree   Potential() (kd-tree monopole, theta 0.5, bucket 8) of the largest 3D group, member order = index order
  pos_out    positions after the type-5 run (AdjustStructureForPeriod applied)

core/core_s3.npz (linking step): KDTree::FOFCriterion(FOF6d, ..., FOFcheckbg) on a halo with two cores in its CM
frame, three iterations of shrinking linking lengths: inputs pos/vel/pot, per iteration p6, p7, minsize and the canonical ids.

veldens/vd_s9.npz (oracle side): GetVelocityDensity on one structure of 6000 float32-exact particles (a halo with two
cores; a fifth of the particles of type 0, which STRUCDEN builds skip): densities of the approximative (per-leaf) and of the exact
(per-particle) neighbour search, Nsearch 256 / Nvel 32 / bucket 16, and of a periodic box with Nsearch 64 / Nvel 8 / bucket 8.

binding/be_s5.npz: GetBindingEnergy on a group-ordered set holding one group above ompunbindnum (tree branch)
and a handful below (pair loop): inputs pos/vel/mass/num/gcmvel, reference outputs phi, E, T, Pot, Efrac, iminpot, ibound.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "velociraptor-stf_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import vroracle  # noqa: E402
import vrsynth  # noqa: E402

CASES = [("box20_s1", 20, 1), ("box24_s2", 24, 2), ("box32_s7", 32, 7)]


def main():
    ref = vroracle.Ref("serial")
    for name, nside, seed in CASES:
        d = vrsynth.make_box(nside ** 3, seed)
        base = dict(period=vrsynth.BOX, ellxscale=vrsynth.BOX / nside, omp_fof=0)
        r5 = ref.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], vroracle.default_opts(fofbgtype=5, **base))
        r3 = ref.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], vroracle.default_opts(fofbgtype=3, **base))
        ru = ref.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], vroracle.default_opts(fofbgtype=5, bound_halos=1, **base))
        c5, _ = vrsynth.canonical_groups(r5["pfof"], d["pid"])
        c3, _ = vrsynth.canonical_groups(r3["pfof"], d["pid"])
        cu, _ = vrsynth.canonical_groups(ru["pfof"], d["pid"])
        sel = np.flatnonzero(c5 == 1)
        phi = ref.potential(r5["pos"][sel], d["mass"][sel], vroracle.default_opts())
        np.savez_compressed(os.path.join(HERE, name + ".npz"), nside=nside, seed=seed, box=vrsynth.BOX,
                            pos=d["pos"].astype(np.float32), vel=d["vel"].astype(np.float32),
                            mass=np.float32(d["mass"][0]), pfof3d=c5.astype(np.int32), pfof6d=c3.astype(np.int32),
                            pfof_unb=cu.astype(np.int32), phi_tree=phi, pos_out=r5["pos"],
                            ng3d=int(c5.max()), ng6d=int(c3.max()), ngunb=int(cu.max()))
        print(name, "groups 3d/6d/unbound:", int(c5.max()), int(c3.max()), int(cu.max()), "largest", len(sel))


def core_search_case():
    ref = vroracle.Ref("serial")
    n, seed = 12000, 3
    pos, vel, pot = vrsynth.halo_with_cores(n, seed)
    pid = np.arange(n)
    p6, p7 = vrsynth.core_search_lengths(pos, vel)
    out = dict(pos=pos.astype(np.float32), vel=vel.astype(np.float32), pot=pot, p9=0.5, minsize=max(20, int(n * 0.005)))
    for it in range(3):
        pr, nr = ref.fofcriterion6d(pos, vel, pot, p6, p7, 0.5, out["minsize"], 16)
        cr, _ = vrsynth.canonical_groups(pr, pid)
        out[f"p6_{it}"] = p6; out[f"p7_{it}"] = p7; out[f"pfof_{it}"] = cr.astype(np.int32); out[f"ng_{it}"] = nr
        print("core search iteration", it, "groups", nr, "largest", int(np.bincount(cr)[1:].max()) if nr else 0)
        p6 *= 0.75 ** 2; p7 *= 0.75 ** 2
    os.makedirs(os.path.join(HERE, "core"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "core", "core_s3.npz"), **out)


def velocity_density_case():
    ref = vroracle.Ref("serial")
    n, seed = 6000, 9
    pos, vel, _ = vrsynth.halo_with_cores(n, seed)
    pos = pos.astype(np.float32).astype(np.float64); vel = vel.astype(np.float32).astype(np.float64)
    rng = np.random.default_rng(seed)
    ptype = (rng.random(n) < 0.8).astype(np.int32)
    out = dict(pos=pos.astype(np.float32), vel=vel.astype(np.float32), ptype=ptype.astype(np.int8))
    out["den_approx"] = ref.velocity_density(pos, vel, 0.0, 16, 256, 32, 1, ptype)
    out["den_exact"] = ref.velocity_density(pos, vel, 0.0, 16, 256, 32, 0, ptype)
    box = np.mod(pos * 40.0, 100.0).astype(np.float32).astype(np.float64)
    out["box"] = box.astype(np.float32)
    out["den_box_approx"] = ref.velocity_density(box, vel, 100.0, 8, 64, 8, 1, None)
    out["den_box_exact"] = ref.velocity_density(box, vel, 100.0, 8, 64, 8, 0, None)
    os.makedirs(os.path.join(HERE, "veldens"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "veldens", "vd_s9.npz"), **out)
    print("vd_s9", "typed", int(ptype.sum()), "median density", float(np.median(out["den_approx"][ptype > 0])))


def binding_energy_case():
    ref = vroracle.Ref("serial")
    rng = np.random.default_rng(5)
    sizes = [1400, 600, 151, 150, 40, 20]
    pos, vel = [], []
    for k, s in enumerate(sizes):
        c = rng.random(3) * 80 + 10
        r = rng.random(s) ** 1.5 * (0.2 + 0.1 * k)
        u = rng.normal(size=(s, 3))
        pos.append(c + r[:, None] * u / np.linalg.norm(u, axis=1)[:, None])
        vel.append(rng.normal(0, 300, 3) + rng.normal(size=(s, 3)) * (60 + 20 * k))
    pos = np.concatenate(pos); vel = np.concatenate(vel).astype(np.float32).astype(np.float64)
    mass = (0.02 * (1 + (rng.random(len(pos)) < 0.2))).astype(np.float32).astype(np.float64)
    num = np.array([0] + sizes, dtype=np.int64)
    start = np.concatenate([[0, 0], np.cumsum(sizes)])
    gcm = np.zeros((len(sizes) + 1, 3))
    for g in range(1, len(sizes) + 1):
        sl = slice(start[g], start[g + 1])
        gcm[g] = (vel[sl] * mass[sl, None]).sum(0) / mass[sl].sum()
    r = ref.binding_energy(pos, vel, mass, num, gcm, vroracle.default_opts())
    os.makedirs(os.path.join(HERE, "binding"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "binding", "be_s5.npz"), pos=pos, vel=vel, mass=mass, num=num, gcmvel=gcm, **r)
    print("be_s5", "groups", len(sizes), "Efrac", np.round(r["Efrac"][1:], 3))


if __name__ == "__main__":
    if "core" in sys.argv[1:]:
        core_search_case()
    elif "veldens" in sys.argv[1:]:
        velocity_density_case()
    else:
        core_search_case()
        velocity_density_case()
        binding_energy_case()
        main()
