"""Timing of vrgpu_binding_energy on the FOF groups of a synthetic box: host buffers in, per-particle
phi / E and per-group figures out (the e2e number of this row), plus the per-stage device times.
usage: python tools/bench_energy.py [nside=128] [repeats=3]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "velociraptor-stf_b200"))
import vrgpu  # noqa: E402
import vrsynth  # noqa: E402

nside = int(sys.argv[1]) if len(sys.argv) > 1 else 128
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
d = vrsynth.make_box(nside ** 3, 3)
ctx = vrgpu.Context(0)
ell = vrsynth.BOX / nside
parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
pfof, ng = ctx.search_full_set(parts, vrgpu.search_opts(vrsynth.BOX, (ell * ell) * (0.2 * 0.2), halo_min_size=20, fofbgtype=5))
order = np.argsort(pfof, kind="stable")
order = order[pfof[order] > 0]
num = np.bincount(pfof, minlength=ng + 1)[1:]
off = np.concatenate([[0], np.cumsum(num)])
gp = np.ascontiguousarray(parts[order])
mv = gp["vel"] * gp["mass"][:, None]
gcm = np.add.reduceat(mv, off[:-1], axis=0) / np.add.reduceat(gp["mass"], off[:-1])[:, None]
G = 43.0211349
ctx.binding_energy(gp, off, gcm, G, 0.0, 0.7, 8, 1000)
ctx.kernel_times_reset()
t0 = time.perf_counter()
for _ in range(reps):
    r = ctx.binding_energy(gp, off, gcm, G, 0.0, 0.7, 8, 1000)
dt = (time.perf_counter() - t0) / reps
kt = ctx.kernel_times()
print(json.dumps({"row": "f1 GetBindingEnergy", "nside": nside, "grouped_particles": int(off[-1]), "groups": int(ng),
                  "largest_group": int(num.max()), "ms_per_call_e2e": dt * 1e3, "Mpart_per_s_e2e": off[-1] / dt / 1e6,
                  "mean_Efrac": float(r["Efrac"].mean()), "stages_ms_per_call": {k: v["ms"] / reps for k, v in kt.items()}}))
