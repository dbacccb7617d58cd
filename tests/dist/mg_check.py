"""Multi-GPU parity check, launched as `torchrun --nproc-per-node N tests/dist/mg_check.py [nside]` on a box with N GPUs.

Every rank builds the same synthetic box, keeps its x-slab, runs vrgpu_dist_search_full_set; rank 0 also runs the
single-GPU path on the whole box and compares the per-particle GLOBAL group ids (must be identical for 3DFOF; for the
6DFOF / unbinding configurations differences are allowed only for marginal particles, whose count is reported)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "velociraptor-stf_b200"))
import vrgpu  # noqa: E402
import vrsynth  # noqa: E402


def main():
    nside = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    ctx = vrgpu.Context(local)
    uid = [vrgpu.dist_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.dist_init(world, rank, uid[0])
    d = vrsynth.make_box(nside ** 3, 40 + nside)
    L = vrsynth.BOX
    lo, hi = L * rank / world, L * (rank + 1) / world
    x = d["pos"][:, 0]
    sel = np.flatnonzero((x >= lo) & (x < hi))
    parts = vrgpu.particles_from_arrays(d["pos"][sel], d["vel"][sel], d["mass"][sel], d["pid"][sel])
    ell = L / nside
    param1 = (ell * ell) * (0.2 * 0.2) * (1.0 * 1.0)
    configs = [("3dfof", dict(fofbgtype=5, bound_halos=0)),
               ("3dfof+unbind(direct)", dict(fofbgtype=5, bound_halos=1, unbind=vrgpu.default_unbind_opts(pp_threshold=-1))),
               ("3dfof+unbind(tree)", dict(fofbgtype=5, bound_halos=1, unbind=vrgpu.default_unbind_opts(pp_threshold=150))),
               ("6dfof+unbind(direct)", dict(fofbgtype=3, bound_halos=1, unbind=vrgpu.default_unbind_opts(pp_threshold=-1)))]
    ok = True
    for name, kw in configs:
        so = vrgpu.search_opts(L, param1, halo_min_size=20, **kw)
        pf, ng = ctx.dist_search_full_set(parts, lo, hi, so)
        gathered = [None] * world
        dist.all_gather_object(gathered, (sel, pf, ng))
        if rank == 0:
            full = np.zeros(nside ** 3, dtype=np.int64)
            for s, p, g in gathered:
                full[s] = p
                assert g == ng
            ref_ctx = vrgpu.Context(local)
            allp = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
            want, ngw = ref_ctx.search_full_set(allp, so)
            nmarg = ref_ctx.stats()["n_marginal"]
            ref_ctx.close()
            ndiff = int(np.count_nonzero(full != want))
            nbound = int(np.count_nonzero((full > 0) != (want > 0)))
            exact = name == "3dfof"
            good = (ndiff == 0 and ng == ngw) if exact else (nbound <= 2 * nmarg + 2)
            ok = ok and good
            print(f"[mg_check] world={world} nside={nside} {name}: groups dist={ng} single={ngw} label diffs={ndiff} "
                  f"bound-set diffs={nbound} marginal={nmarg} -> {'OK' if good else 'FAIL'}", flush=True)
    res = [ok]
    dist.broadcast_object_list(res, src=0)
    ctx.close()
    dist.destroy_process_group()
    sys.exit(0 if res[0] else 1)


if __name__ == "__main__":
    main()
