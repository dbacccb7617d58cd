"""One-GPU probe: time the direct-sum kernels on the group subset one rank of an N-rank run would own (LPT placement)."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "velociraptor-stf_b200"))
import vrgpu, vrsynth

nside = 256
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
d = vrsynth.make_box(nside ** 3, 102)
ctx = vrgpu.Context(0)
parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
b = 0.2 * vrsynth.BOX / nside
pf, ng = ctx.fof3d(parts, b, vrsynth.BOX, 20)
sizes = np.bincount(pf, minlength=ng + 1)[1:].astype(float)
first = np.full(ng + 1, -1, dtype=np.int64)
idx = np.arange(len(pf)); m = pf > 0
first[pf[m][::-1]] = idx[m][::-1]
home = np.minimum((d["pos"][first[1:], 0] / (vrsynth.BOX / N)).astype(int), N - 1)
cost = lambda n: 0.55e-3 * n * n
load = np.zeros(N)
small = sizes < 2048
for r in range(N):
    load[r] = cost(sizes[small & (home == r)]).sum()
dest = home.copy()
for g in np.flatnonzero(~small):
    r = int(np.argmin(load)); load[r] += cost(sizes[g]); dest[g] = r
print("loads", np.round(load / load.max(), 3))
uo = vrgpu.default_unbind_opts(pp_threshold=-1)
for label, keep in (("rank0 all", dest == 0), ("rank0 >=4096", (dest == 0) & (sizes >= 4096)), ("rank0 <4096", (dest == 0) & (sizes < 4096)),
                    ("all groups", np.ones(ng, bool)), ("all <4096", sizes < 4096)):
    remap = np.zeros(ng + 1, dtype=np.int64)
    remap[1:][keep] = np.arange(1, int(keep.sum()) + 1)
    sub = remap[pf]
    for rep in range(2):
        ctx.kernel_times_reset()
        ctx.unbind(parts, sub, int(keep.sum()), uo)
        kt = ctx.kernel_times()
    print(label, "groups", int(keep.sum()), "members", int((sub > 0).sum()),
          {k: round(v["ms"], 2) for k, v in kt.items() if "potential" in k})
