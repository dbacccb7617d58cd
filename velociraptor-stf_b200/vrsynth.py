"""Synthetic clustered periodic boxes.

Periodic cube L=100 (code length, Mpc), N particles; Omega_m=0.3, h=1, a=1 so that the particle mass is
m = Omega_m * rho_crit * L^3 / N with rho_crit = 27.755 (1e10 Msun / Mpc^3) and G = 43.0211349 as in the
reference's examples/sample_dmcosmological_run.cfg.  Half of the particles sit in N/4000 haloes whose sizes
follow dN/dM ~ M^-1.9 between 20 and N/50 particles with an NFW (c=5) radial profile truncated at R200; the
other half is uniform.  Velocities: halo bulk N(0,300 km/s) + isotropic Gaussian sigma_v = sqrt(G M / 2 R200);
background N(0,50).  Everything is rounded to float32 (exactly what the reference's Gadget reader promotes to
double, gadgetio.cxx:258,770), then returned as float64 arrays.
"""
from __future__ import annotations

import numpy as np

G_CODE = 43.0211349
RHO_CRIT = 27.755
OMEGA_M = 0.3
BOX = 100.0


def particle_mass(n: int, box: float = BOX) -> float:
    return OMEGA_M * RHO_CRIT * box ** 3 / n


def linking_length(n: int, b: float = 0.2, box: float = BOX) -> float:
    """Physical 3DFOF linking length: ellxscale*ellphys*ellhalophysfac with ellxscale = LN = L / N^(1/3)
    (gadgetio.cxx:219,1417; search.cxx:77)."""
    return b * box / round(n ** (1.0 / 3.0)) if round(n ** (1.0 / 3.0)) ** 3 == n else b * box / n ** (1.0 / 3.0)


def _nfw_inverse_table(c: float, npts: int = 4096):
    x = np.linspace(0.0, c, npts)
    m = np.log1p(x) - x / (1.0 + x)
    m /= m[-1]
    return m, x / c  # cumulative mass fraction -> r / R200


def halo_sizes(n: int, rng: np.random.Generator, frac: float = 0.5, nmin: int = 20, slope: float = -1.9):
    # SURVEY §8(d) asks for N/4000 haloes AND a M^-1.9 spectrum on [20, N/50] AND 50 % of the mass in haloes;
    # the three are mutually inconsistent (the spectrum's mean is ~300, not 2000), so the halo COUNT is the
    # free parameter here: sizes are drawn from the spectrum until they hold `frac` of the particles.
    nmax = max(nmin + 1, n // 50)
    a = slope + 1.0
    target = int(frac * n)
    chunks, tot = [], 0
    while tot < target:
        u = rng.random(max(16, target // 200))
        s = np.floor((nmin ** a + u * (nmax ** a - nmin ** a)) ** (1.0 / a)).astype(np.int64)
        chunks.append(s)
        tot += int(s.sum())
    s = np.concatenate(chunks)
    k = int(np.searchsorted(np.cumsum(s), target, side="left")) + 1
    s = s[:k]
    s[-1] -= int(s.sum()) - target
    if s[-1] < nmin:  # fold a too-small remainder into the previous halo
        if k > 1:
            s[-2] += s[-1]
            s = s[:-1]
    return np.sort(s)[::-1]


def make_box(n: int, seed: int, box: float = BOX, conc: float = 5.0, frac: float = 0.5,
             single_halo: bool = False):
    """Return dict(pos[n,3], vel[n,3], mass[n], pid[n]) as float64 / int64, float32-exact values.

    single_halo=True builds the config-5 style input: one NFW halo (c=7) holding 90 % of the particles plus
    10 % in subclumps, placed in the box centre."""
    rng = np.random.default_rng(seed)
    m = particle_mass(n, box)
    pos = np.empty((n, 3), dtype=np.float64)
    vel = np.empty((n, 3), dtype=np.float64)
    if single_halo:
        nsub_tot = n // 10
        nsub = max(1, min(2000, nsub_tot // 50))
        sizes = np.concatenate([[n - nsub_tot], np.full(nsub, nsub_tot // nsub, dtype=np.int64)])
        sizes[0] += n - int(sizes.sum())
        conc_arr = np.concatenate([[7.0], np.full(nsub, 5.0)])
        nhalo_part = n
    else:
        sizes = halo_sizes(n, rng, frac)
        conc_arr = np.full(len(sizes), conc)
        nhalo_part = int(sizes.sum())
    nh = len(sizes)
    mass_h = sizes * m
    r200 = (3.0 * mass_h / (800.0 * np.pi * RHO_CRIT)) ** (1.0 / 3.0)
    if single_halo:
        centres = np.empty((nh, 3))
        centres[0] = 0.5 * box
        # subclumps distributed inside the main halo following its profile
        mt, rt = _nfw_inverse_table(7.0)
        rr = np.interp(rng.random(nh - 1), mt, rt) * r200[0]
        d = rng.normal(size=(nh - 1, 3))
        d /= np.linalg.norm(d, axis=1)[:, None]
        centres[1:] = centres[0] + rr[:, None] * d
        bulk = np.zeros((nh, 3))
        sig0 = np.sqrt(G_CODE * mass_h[0] / (2.0 * r200[0]))
        bulk[1:] = rng.normal(0.0, sig0, size=(nh - 1, 3))
    else:
        centres = rng.random((nh, 3)) * box
        bulk = rng.normal(0.0, 300.0, size=(nh, 3))
    sigv = np.sqrt(G_CODE * mass_h / (2.0 * r200))
    hid = np.repeat(np.arange(nh), sizes)
    tables = {}
    u = rng.random(nhalo_part)
    rfrac = np.empty(nhalo_part)
    for c in np.unique(conc_arr):
        if c not in tables:
            tables[c] = _nfw_inverse_table(float(c))
        sel = conc_arr[hid] == c
        rfrac[sel] = np.interp(u[sel], tables[c][0], tables[c][1])
    d = rng.normal(size=(nhalo_part, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    pos[:nhalo_part] = centres[hid] + (rfrac * r200[hid])[:, None] * d
    vel[:nhalo_part] = bulk[hid] + rng.normal(size=(nhalo_part, 3)) * sigv[hid][:, None]
    nbg = n - nhalo_part
    if nbg > 0:
        pos[nhalo_part:] = rng.random((nbg, 3)) * box
        vel[nhalo_part:] = rng.normal(0.0, 50.0, size=(nbg, 3))
    # shuffle so that input order carries no group information
    perm = rng.permutation(n)
    pos = pos[perm]
    vel = vel[perm]
    pos = np.mod(pos, box)
    pos = pos.astype(np.float32).astype(np.float64)
    pos[pos >= box] = np.float64(np.nextafter(np.float32(box), np.float32(0)))
    vel = vel.astype(np.float32).astype(np.float64)
    mass = np.full(n, np.float64(np.float32(m)))
    pid = np.arange(1, n + 1, dtype=np.int64)
    return {"pos": np.ascontiguousarray(pos), "vel": np.ascontiguousarray(vel), "mass": mass, "pid": pid,
            "box": box, "n": n, "seed": seed}


def canonical_groups(pfof: np.ndarray, pid: np.ndarray | None = None):
    """Canonical relabelling used for parity (BASELINE.json): groups ordered by size descending,
    ties by minimum particle PID ascending; returns (labels 0..G, sizes[G])."""
    pfof = np.asarray(pfof)
    n = pfof.shape[0]
    if pid is None:
        pid = np.arange(n, dtype=np.int64)
    g = pfof > 0
    if not g.any():
        return np.zeros(n, dtype=np.int64), np.zeros(0, dtype=np.int64)
    ids, inv, cnt = np.unique(pfof[g], return_inverse=True, return_counts=True)
    minpid = np.full(len(ids), np.iinfo(np.int64).max)
    np.minimum.at(minpid, inv, pid[g])
    order = np.lexsort((minpid, -cnt))
    rank = np.empty(len(ids), dtype=np.int64)
    rank[order] = np.arange(1, len(ids) + 1)
    out = np.zeros(n, dtype=np.int64)
    out[g] = rank[inv]
    return out, cnt[order]


# ------------------------------------------------------------------------------------------------------
# slab-wise generator: the same clustered box for ANY partition into x-slabs (multi-GPU runs at 512^3 / 1024^3, where no
# rank can hold the whole box).  Every halo and every background bin has its own random stream, so a rank generates only
# what can reach its slab.
# ------------------------------------------------------------------------------------------------------
SLAB_BINS = 64          # background is drawn per x-bin of width box/64: slab faces must be multiples of that


def _scramble_pid(i: np.ndarray, n: int) -> np.ndarray:
    """Bijection of [0, n) (n a power of two times an odd number is fine: the multiplier is coprime with n)."""
    a = 2654435761
    while np.gcd(a, n) != 1:
        a += 2
    return (i.astype(np.uint64) * np.uint64(a) + np.uint64(12345)) % np.uint64(n)


def make_slab(n: int, seed: int, lo: float, hi: float, box: float = BOX, conc: float = 5.0, frac: float = 0.5):
    """Particles of the n-particle clustered box whose x lies in [lo, hi); same statistics as make_box (NOT the same
    realisation).  The union over any set of slabs tiling [0, box) with faces on multiples of box/64 is one fixed particle
    set.  pid is a scrambled global index (1..n) and the particles come sorted by pid, so array order == pid order on
    every rank and for the whole box alike."""
    rng0 = np.random.default_rng(seed)
    m = particle_mass(n, box)
    sizes = halo_sizes(n, rng0, frac)
    nh = len(sizes)
    nhp = int(sizes.sum())
    centres = rng0.random((nh, 3)) * box
    bulk = rng0.normal(0.0, 300.0, size=(nh, 3))
    mass_h = sizes * m
    r200 = (3.0 * mass_h / (800.0 * np.pi * RHO_CRIT)) ** (1.0 / 3.0)
    sigv = np.sqrt(G_CODE * mass_h / (2.0 * r200))
    first = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64)
    mt, rt = _nfw_inverse_table(conc)
    # halos that can reach [lo, hi) (periodic in x); float32 rounding moves a particle by < 1e-5
    cx = centres[:, 0]
    pad = r200 + 1e-3
    inside = (cx >= lo) & (cx < hi)
    dlo = np.abs(cx - lo); dlo = np.minimum(dlo, box - dlo)
    dhi = np.abs(cx - hi); dhi = np.minimum(dhi, box - dhi)
    sel = np.flatnonzero(inside | (np.minimum(dlo, dhi) <= pad))
    P, V, I = [], [], []
    top = np.float64(np.nextafter(np.float32(box), np.float32(0)))

    def keep(pos, vel, idx):
        pos = np.mod(pos, box).astype(np.float32).astype(np.float64)
        pos[pos >= box] = top
        k = (pos[:, 0] >= lo) & (pos[:, 0] < hi)
        if k.any():
            P.append(pos[k]); V.append(vel[k].astype(np.float32).astype(np.float64)); I.append(idx[k])

    for h in sel:
        s = int(sizes[h])
        rng = np.random.default_rng([seed, 1, int(h)])
        u = rng.random(s)
        d = rng.normal(size=(s, 3))
        w = rng.normal(size=(s, 3))
        d /= np.linalg.norm(d, axis=1)[:, None]
        pos = centres[h] + (np.interp(u, mt, rt) * r200[h])[:, None] * d
        keep(pos, bulk[h] + w * sigv[h], first[h] + np.arange(s, dtype=np.int64))
    nbg = n - nhp
    per = nbg // SLAB_BINS
    extra = nbg - per * SLAB_BINS
    wbin = box / SLAB_BINS
    b0 = int(np.floor(lo / wbin + 1e-9)) - 1          # one bin further left: float32 rounding can reach the bin's upper edge
    b1 = int(np.ceil(hi / wbin - 1e-9))
    for bb in sorted({b % SLAB_BINS for b in range(b0, b1)}):
        cnt = per + (1 if bb < extra else 0)
        if cnt == 0:
            continue
        rng = np.random.default_rng([seed, 2, bb])
        pos = rng.random((cnt, 3)) * np.array([wbin, box, box]) + np.array([bb * wbin, 0.0, 0.0])
        vel = rng.normal(0.0, 50.0, size=(cnt, 3))
        start = nhp + bb * per + min(bb, extra)
        keep(pos, vel, start + np.arange(cnt, dtype=np.int64))
    if P:
        pos = np.concatenate(P); vel = np.concatenate(V); idx = np.concatenate(I)
    else:
        pos = np.zeros((0, 3)); vel = np.zeros((0, 3)); idx = np.zeros(0, dtype=np.int64)
    pid = _scramble_pid(idx, n).astype(np.int64) + 1
    o = np.argsort(pid, kind="stable")
    nn = len(pid)
    return {"pos": np.ascontiguousarray(pos[o]), "vel": np.ascontiguousarray(vel[o]), "mass": np.full(nn, np.float64(np.float32(m))),
            "pid": np.ascontiguousarray(pid[o]), "box": box, "n": nn, "seed": seed, "n_total": n}


def halo_with_cores(n, seed):
    """One halo in its CM frame with two dense, cold cores (a merger remnant) and a few already-tagged substructure
    particles -- the input of the 6D core search (search.cxx:1531-1700)."""
    rng = np.random.default_rng(seed)
    n1, n2 = n // 6, n // 10
    nb = n - n1 - n2
    r = 0.8 * rng.random(nb) ** 1.2
    d = rng.normal(size=(nb, 3)); d /= np.linalg.norm(d, axis=1)[:, None]
    pos = [r[:, None] * d, np.array([0.25, 0.0, 0.0]) + 0.03 * rng.normal(size=(n1, 3)),
           np.array([-0.3, 0.1, 0.0]) + 0.025 * rng.normal(size=(n2, 3))]
    vel = [250.0 * rng.normal(size=(nb, 3)), np.array([150.0, 0, 0]) + 40.0 * rng.normal(size=(n1, 3)),
           np.array([-200.0, 50.0, 0]) + 35.0 * rng.normal(size=(n2, 3))]
    pos = np.concatenate(pos); vel = np.concatenate(vel)
    pos = pos.astype(np.float32).astype(np.float64); vel = vel.astype(np.float32).astype(np.float64)
    tagged = rng.random(n) < 0.03                      # pfof > 0: Partsubset[i].SetPotential(pfof) (search.cxx:1605)
    pot = np.where(tagged, rng.integers(1, 5, size=n).astype(np.float64), 0.0)
    return pos, vel, pot


def core_search_lengths(pos, vel):
    """params[6], params[7] of the first 6D core search of such a halo: a fraction of the mean inter-particle spacing and
    of the velocity dispersion, in the manner of search.cxx:1586-1598."""
    n = len(pos)
    ell = 0.2 * (4.0 / 3.0 * np.pi * 0.8 ** 3 / n) ** (1.0 / 3.0) * 3.0
    sig2 = float(np.mean(np.sum((vel - vel.mean(axis=0)) ** 2, axis=1)) / 3.0)
    return (ell * 0.6) ** 2, sig2 * 0.5 ** 2
