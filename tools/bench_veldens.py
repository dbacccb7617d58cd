"""Timing of vrgpu_velocity_density on one synthetic structure: host buffers in, densities out (the e2e
number of this row), the per-stage device times, and -- with --cpu -- the reference's own GetVelocityDensity (OpenMP build
of oracle/_ref) on a bounded sample of the same structure, checked against the device result on that sample.
usage: python tools/bench_veldens.py [n=1000000] [repeats=3] [--exact] [--cpu]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "velociraptor-stf_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import vrgpu  # noqa: E402
import vrsynth  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
n = int(args[0]) if len(args) > 0 else 1000000
reps = int(args[1]) if len(args) > 1 else 3
approx = 0 if "--exact" in sys.argv else 1
pos, vel, _ = vrsynth.halo_with_cores(n, 9)
ctx = vrgpu.Context(0)
parts = vrgpu.particles_from_arrays(pos, vel)
den = ctx.velocity_density(parts, approx=approx)
ctx.kernel_times_reset()
t0 = time.perf_counter()
for _ in range(reps):
    den = ctx.velocity_density(parts, approx=approx)
dt = (time.perf_counter() - t0) / reps
kt = ctx.kernel_times()
out = {"row": "f4 GetVelocityDensity", "n": n, "approx": approx, "nsearch": 256, "nvel": 32, "bsize": 16, "ms_per_call_e2e": dt * 1e3,
       "Mpart_per_s_e2e": n / dt / 1e6, "median_density": float(np.median(den)),
       "stages_ms_per_call": {k: v["ms"] / reps for k, v in kt.items()}}
if "--cpu" in sys.argv:
    import vroracle
    m = min(n, 200000 if approx else 50000)        # a structure of its own: the first m particles
    ref = vroracle.Ref("omp" if vroracle.Ref.available("omp") else "serial")
    t0 = time.perf_counter()
    want = ref.velocity_density(pos[:m], vel[:m], approx=approx)
    tc = time.perf_counter() - t0
    got = ctx.velocity_density(vrgpu.particles_from_arrays(pos[:m], vel[:m]), approx=approx)
    out["cpu_reference"] = {"sample": f"first {m} particles as one structure", "seconds": tc, "Mpart_per_s": m / tc / 1e6,
                            "cores": os.cpu_count(), "max_rel_diff_vs_device": float(np.max(np.abs(got / want - 1.0)))}
print(json.dumps(out))
