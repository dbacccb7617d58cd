"""CPU model of the multi-GPU protocol of velociraptor-stf_b200/csrc/multigpu.cu, run with torch.distributed (gloo).

Same steps as the device code -- x-slabs, ghosts within b of a face (periodic wrap applied to the ghost), local linking of
owned + ghost particles, one edge per ghost (my component, the owner's component), all-gather of the edges, replicated
union-find over the boundary components, sizes summed over ranks, all-gathered candidate records ordered by
(size desc, min pid asc) -- with the oracle's open-box FOF standing in for the local CUDA linker (y/z periodicity is
modelled with image ghosts of the rank's own particles, which yield edges exactly like remote ghosts).
Launched by tests/test_dist_model.py; exits 0 iff the global catalogue equals the single-domain oracle."""
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "velociraptor-stf_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import vroracle  # noqa: E402
import vrsynth  # noqa: E402


def find(parent, a):
    while parent[a] != a:
        parent[a] = parent[parent[a]]
        a = parent[a]
    return a


def main():
    nside = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    orc = vroracle.Oracle()
    d = vrsynth.make_box(nside ** 3, 77)
    L = vrsynth.BOX
    b = 0.2 * L / nside
    bm = b * (1 + 1e-9)
    minsize = 20
    lo, hi = L * rank / world, L * (rank + 1) / world
    x = d["pos"][:, 0]
    own = np.flatnonzero((x >= lo) & (x < hi))
    pos, pid = d["pos"][own], d["pid"][own]
    n_own = len(own)
    left, right = (rank - 1) % world, (rank + 1) % world
    # ---- ghosts to the neighbours (wrap applied by the sender) ----
    toL = np.flatnonzero(pos[:, 0] < lo + bm)
    toR = np.flatnonzero(pos[:, 0] >= hi - bm)
    gL = pos[toL].copy(); gR = pos[toR].copy()
    if rank == 0: gL[:, 0] = gL[:, 0] + L
    if rank == world - 1: gR[:, 0] = gR[:, 0] - L
    box = [None] * world
    dist.all_gather_object(box, {"toL": gL, "toR": gR})
    fromR, fromL = box[right]["toL"], box[left]["toR"]
    ext = np.concatenate([pos, fromR, fromL])
    n_x = len(ext)
    # ---- y/z images of everything held (owned or ghost): origin index recorded ----
    img_pos, img_org = [], []
    for sy in (-1, 0, 1):
        for sz in (-1, 0, 1):
            if sy == 0 and sz == 0:
                continue
            m = np.ones(n_x, dtype=bool)
            if sy == 1: m &= ext[:, 1] < bm
            if sy == -1: m &= ext[:, 1] >= L - bm
            if sz == 1: m &= ext[:, 2] < bm
            if sz == -1: m &= ext[:, 2] >= L - bm
            idx = np.flatnonzero(m)
            p = ext[idx].copy()
            p[:, 1] = p[:, 1] + sy * L
            p[:, 2] = p[:, 2] + sz * L
            img_pos.append(p); img_org.append(idx)
    allp = np.concatenate([ext] + img_pos)
    img_org = np.concatenate(img_org) if img_org else np.zeros(0, dtype=np.int64)
    # ---- local linking (open box), every particle labelled ----
    lab, _ = orc.fof3d(allp, b, None, 1, None)
    comp = lab.astype(np.int64)  # 1..ncomp, every particle in a component (minsize 1)
    # ---- label exchange for the x-ghosts ----
    lbox = [None] * world
    dist.all_gather_object(lbox, {"toL": (rank << 32) | comp[toL], "toR": (rank << 32) | comp[toR]})
    rlab = np.concatenate([lbox[right]["toL"], lbox[left]["toR"]])
    edges = [np.stack([(rank << 32) | comp[n_own:n_x], rlab], axis=1)]
    # images: (component of the image, component of its origin on THIS rank)
    edges.append(np.stack([(rank << 32) | comp[n_x:], (rank << 32) | comp[img_org]], axis=1))
    edges = np.concatenate(edges)
    ebox = [None] * world
    dist.all_gather_object(ebox, edges)
    alle = np.concatenate(ebox)
    # ---- replicated union-find over the boundary components ----
    U, inv = np.unique(alle.reshape(-1), return_inverse=True)
    inv = inv.reshape(-1, 2)
    parent = np.arange(len(U))
    for a, c in inv:
        ra, rc = find(parent, a), find(parent, c)
        if ra != rc:
            parent[max(ra, rc)] = min(ra, rc)
    R = np.array([find(parent, i) for i in range(len(U))], dtype=np.int64)
    # ---- sizes / min pid of OWNED particles per local component ----
    ncomp = int(comp.max())
    csize = np.bincount(comp[:n_own], minlength=ncomp + 1)
    cmin = np.full(ncomp + 1, np.iinfo(np.int64).max)
    np.minimum.at(cmin, comp[:n_own], pid)
    mine = (U >> 32) == rank
    usize = np.where(mine, csize[np.where(mine, U & 0xffffffff, 0)], 0)
    umin = np.where(mine, cmin[np.where(mine, U & 0xffffffff, 0)], np.iinfo(np.int64).max)
    sbox = [None] * world
    dist.all_gather_object(sbox, (usize, umin))
    usize = np.sum([s[0] for s in sbox], axis=0)
    umin = np.min([s[1] for s in sbox], axis=0)
    gsz = np.zeros(len(U), dtype=np.int64); np.add.at(gsz, R, usize)
    gmp = np.full(len(U), np.iinfo(np.int64).max); np.minimum.at(gmp, R, umin)
    bidx = np.full(ncomp + 1, -1, dtype=np.int64)
    bidx[U[mine] & 0xffffffff] = np.flatnonzero(mine)
    # ---- candidate records: interior components of this rank + boundary roots whose label is this rank's ----
    rec = [(int(csize[c]), int(cmin[c]), 0, c) for c in range(1, ncomp + 1) if bidx[c] < 0 and csize[c] >= minsize]
    rec += [(int(gsz[u]), int(gmp[u]), 1, int(u)) for u in range(len(U)) if R[u] == u and (U[u] >> 32) == rank and gsz[u] >= minsize]
    rbox = [None] * world
    dist.all_gather_object(rbox, rec)
    allrec = [(s, m, k, ref, r) for r, lst in enumerate(rbox) for (s, m, k, ref) in lst]
    order = sorted(range(len(allrec)), key=lambda j: (-allrec[j][0], allrec[j][1]))
    gid_u, gid_c = {}, {}
    for g, j in enumerate(order, start=1):
        s, m, k, ref, r = allrec[j]
        if k == 1: gid_u[ref] = g
        elif r == rank: gid_c[ref] = g
    pf = np.zeros(n_own, dtype=np.int64)
    for i in range(n_own):
        c = comp[i]
        pf[i] = gid_u.get(int(R[bidx[c]]), 0) if bidx[c] >= 0 else gid_c.get(int(c), 0)
    # ---- compare with the single-domain oracle ----
    pbox = [None] * world
    dist.all_gather_object(pbox, (own, pf))
    ok = True
    want, ngw = orc.fof3d(d["pos"], b, L, minsize, d["pid"])
    if rank == 0:
        full = np.zeros(nside ** 3, dtype=np.int64)
        for o, p in pbox:
            full[o] = p
        ok = bool(np.array_equal(full, want)) and len(allrec) == ngw
        print(f"[dist_model] world={world} nside={nside} groups={len(allrec)} oracle={ngw} boundary comps={len(U)} edges={len(alle)} -> {'OK' if ok else 'FAIL'}")

    # ---- AdjustStructureForPeriod without moving a particle (mg_adjust_global): the reference particle of a group is its member
    #      with the smallest pid; its position reaches every rank through an all-reduce MIN of the pid and an all-reduce SUM of a
    #      table that only the owner of that pid fills (x + 0 is exact); every member is then shifted by whole periods as
    #      search.cxx:871-887 does.  Result: the single-domain oracle's coordinates, bit for bit (pids ascend with the index here). ----
    import torch
    Gall = len(allrec)
    minpid = torch.full((Gall + 1,), np.iinfo(np.int64).max, dtype=torch.int64)
    for i in range(n_own):
        if pf[i] > 0 and pid[i] < minpid[pf[i]]:
            minpid[pf[i]] = int(pid[i])
    dist.all_reduce(minpid, op=dist.ReduceOp.MIN)
    refpos = torch.zeros((Gall + 1, 3), dtype=torch.float64)
    mp = minpid.numpy()
    for i in range(n_own):
        if pf[i] > 0 and pid[i] == mp[pf[i]]:
            refpos[pf[i]] = torch.from_numpy(pos[i].copy())
    dist.all_reduce(refpos, op=dist.ReduceOp.SUM)
    rp = refpos.numpy()
    adj = pos.copy()
    for i in range(n_own):
        g = pf[i]
        if g <= 0:
            continue
        for k in range(3):
            diff = rp[g, k] - adj[i, k]
            if diff < -0.5 * L:
                adj[i, k] = adj[i, k] - L
            elif diff > 0.5 * L:
                adj[i, k] = adj[i, k] + L
    abox = [None] * world
    dist.all_gather_object(abox, (own, adj))
    if rank == 0:
        fullpos = np.array(d["pos"], dtype=np.float64, copy=True)
        for o, a in abox:
            fullpos[o] = a
        wantpos = orc.adjust_for_period(d["pos"], L, want, ngw)
        same_pos = bool(np.array_equal(fullpos, wantpos))
        nmoved = int(np.count_nonzero(np.any(fullpos != d["pos"], axis=1)))
        print(f"[dist_model] period adjustment over {world} ranks: {nmoved} particles moved, coordinates identical to the single-domain oracle: {same_pos} "
              f"-> {'OK' if same_pos else 'FAIL'}")
        ok = ok and same_pos

    # ---- placement of the per-group stages (mg_place_groups): straddling or costly groups to the least loaded rank in
    #      size order, small single-rank groups stay; then consolidation, unbinding per group, global renumbering ----
    G = len(allrec)
    gsize = np.zeros(G, dtype=np.int64); ghome = np.full(G, -1, dtype=np.int64)
    for g, j in enumerate(order):
        s_, m_, k_, ref_, r_ = allrec[j]
        gsize[g] = s_
        ghome[g] = -1 if k_ == 1 else r_
    BIG = 200                                   # scaled down with the box (2048 in the device code)
    cost = lambda n: 0.55e-3 * n * n if n <= 150 else 0.45 * n * np.log2(n + 1.0) + 2.0 * n
    load = np.zeros(world)
    for g in range(G):
        if ghome[g] >= 0 and gsize[g] < BIG:
            load[ghome[g]] += cost(gsize[g])
    dest_of = np.full(G + 1, -1, dtype=np.int64)
    for g in range(G):
        dd = ghome[g]
        if dd < 0 or gsize[g] >= BIG:
            dd = int(np.argmin(load)); load[dd] += cost(gsize[g])
        dest_of[g + 1] = dd
    dest = dest_of[pf]
    vel, mass = d["vel"][own], d["mass"][own]
    exports = [None] * world
    for t in range(world):
        m = (dest == t) & (t != rank)
        exports[t] = (pos[m], vel[m], mass[m], pid[m], pf[m], own[m])
    xbox = [None] * world
    dist.all_gather_object(xbox, exports)
    keep = (dest < 0) | (dest == rank)
    wpos = [pos]; wvel = [vel]; wmass = [mass]; wpid = [pid]; wpf = [np.where(keep, pf, 0)]; wown = [own]
    for s_ in range(world):
        if s_ == rank:
            continue
        e = xbox[s_][rank]
        wpos.append(e[0]); wvel.append(e[1]); wmass.append(e[2]); wpid.append(e[3]); wpf.append(e[4]); wown.append(e[5])
    wpos, wvel, wmass, wpid, wpf, wown = (np.concatenate(a) for a in (wpos, wvel, wmass, wpid, wpf, wown))
    present = np.unique(wpf[wpf > 0])
    local_id = np.zeros(G + 1, dtype=np.int64); local_id[present] = np.arange(1, len(present) + 1)
    wl = local_id[wpf]
    # members in pid order within a group (what by_pid=true does on the device): sort the working set by pid
    o_pid = np.argsort(wpid, kind="stable")
    wpos, wvel, wmass, wpid, wl, wown = wpos[o_pid], wvel[o_pid], wmass[o_pid], wpid[o_pid], wl[o_pid], wown[o_pid]
    uo = vroracle.unbind_opts_from(vroracle.default_opts(), pp_threshold=150, sortflag_mode=0)
    wposu = orc.adjust_for_period(wpos, L, wl, len(present)) if len(present) else wpos
    res_u = orc.unbind(wposu, wvel, wmass, wl, len(present), uo, wpid, 1e-6) if len(present) else {"pfof": wl, "ngroups": 0, "n_marginal": 0}
    fl = res_u["pfof"]
    frec = []
    for g in range(1, int(res_u["ngroups"]) + 1):
        m = fl == g
        frec.append((int(m.sum()), int(wpid[m].min()), g))
    fbox = [None] * world
    dist.all_gather_object(fbox, frec)
    fall = [(s_, m_, g_, r_) for r_, lst in enumerate(fbox) for (s_, m_, g_) in lst]
    forder = sorted(range(len(fall)), key=lambda j: (-fall[j][0], fall[j][1]))
    fgid = {(fall[j][3], fall[j][2]): g for g, j in enumerate(forder, start=1)}
    fglob = np.array([fgid[(rank, int(x))] if x > 0 else 0 for x in fl], dtype=np.int64)
    hbox = [None] * world
    dist.all_gather_object(hbox, (wown, fglob, wl > 0, int(res_u["n_marginal"])))
    if rank == 0:
        full2 = np.zeros(nside ** 3, dtype=np.int64); seen = np.zeros(nside ** 3, dtype=np.int64)
        nmarg = 0
        for o_, f_, held, nm in hbox:
            full2[o_[held]] = f_[held]
            np.add.at(seen, o_[held], 1)
            nmarg += nm
        posw = orc.adjust_for_period(d["pos"], L, want, ngw)
        ref_u = orc.unbind(posw, d["vel"], d["mass"], want, ngw, uo, d["pid"], 1e-6)
        once = bool(np.array_equal(seen, (want > 0).astype(np.int64)))      # every grouped particle consolidated exactly once
        ndiff = int(np.count_nonzero((full2 > 0) != (ref_u["pfof"] > 0)))
        same = ndiff <= nmarg + ref_u["n_marginal"] and (ndiff > 0 or (np.array_equal(full2, ref_u["pfof"]) and len(fall) == ref_u["ngroups"]))
        print(f"[dist_model] placement: loads={np.round(load / max(load.max(), 1e-30), 3).tolist()} consolidated once={once} "
              f"unbound parity diffs={ndiff} (marginal {nmarg}) groups={len(fall)} oracle={ref_u['ngroups']} -> {'OK' if once and same else 'FAIL'}")
        ok = ok and once and same
    res = [ok]
    dist.broadcast_object_list(res, src=0)
    dist.destroy_process_group()
    sys.exit(0 if res[0] else 1)


if __name__ == "__main__":
    main()
