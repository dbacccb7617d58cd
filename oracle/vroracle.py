"""oracle/vroracle.py -- TEST INFRASTRUCTURE ONLY (ctypes loaders for the checkers).

Two checkers live under oracle/:
  * `ref`    : the UNMODIFIED reference compiled into oracle/_ref/libvrref.so (serial, parity truth) and
               oracle/_ref/libvrref_omp.so (OpenMP, timed CPU baseline) -- see oracle/Makefile, ref_driver.cxx.
  * `oracle` : our plain-C restatement oracle/liboracle.so (vr_oracle.c), pinned against `ref`.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module.  The product (velociraptor-stf_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

c_ll = C.c_longlong
c_dbl = C.c_double
P_ll = C.POINTER(C.c_longlong)
P_dbl = C.POINTER(C.c_double)


class RefOpts(C.Structure):
    """Mirror of vrref_opts in oracle/ref_driver.cxx (same order)."""
    _fields_ = [
        ("period", c_dbl), ("ellxscale", c_dbl), ("ellphys", c_dbl), ("ellhalophysfac", c_dbl),
        ("ellhalo6dxfac", c_dbl), ("ellhalo6dvfac", c_dbl),
        ("halo_min_size", c_ll), ("min_size", c_ll),
        ("fofbgtype", C.c_int), ("bound_halos", C.c_int), ("omp_fof", C.c_int),
        ("omp_fof_size", c_ll),
        ("G", c_dbl), ("eps", c_dbl), ("Eratio", c_dbl), ("minEfrac", c_dbl),
        ("unbindtype", C.c_int), ("bgpot", C.c_int), ("cmvelreftype", C.c_int),
        ("fracpotref", c_dbl), ("Npotref", c_ll),
        ("maxunbindfrac", c_dbl), ("maxunboundfracforiterativeunbind", c_dbl), ("maxallowedunboundfrac", c_dbl),
        ("TreeThetaOpen", c_dbl), ("BucketSize", C.c_int), ("verbose", C.c_int),
        ("iterate_cm", C.c_int), ("sort_by_binding_energy", C.c_int),
    ]


def default_opts(**kw) -> RefOpts:
    """Defaults = reference defaults (allvars.h:312-329, 660-847) + the sample cfg's unbinding block."""
    o = RefOpts()
    o.period = 0.0
    o.ellxscale = 1.0
    o.ellphys = 0.2
    o.ellhalophysfac = 1.0
    o.ellhalo6dxfac = 1.0
    o.ellhalo6dvfac = 1.25
    o.halo_min_size = 20
    o.min_size = 20
    o.fofbgtype = 5
    o.bound_halos = 0
    o.omp_fof = 0
    o.omp_fof_size = 0
    o.G = 43.0211349
    o.eps = 0.0
    o.Eratio = 0.95
    o.minEfrac = 0.65
    o.unbindtype = 0
    o.bgpot = 1
    o.cmvelreftype = 0
    o.fracpotref = 1.0
    o.Npotref = 20
    o.maxunbindfrac = 0.5
    o.maxunboundfracforiterativeunbind = 0.95
    o.maxallowedunboundfrac = 0.025
    o.TreeThetaOpen = 0.5
    o.BucketSize = 8
    o.verbose = 0
    o.iterate_cm = 1
    o.sort_by_binding_energy = 1
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _pd(a):
    return a.ctypes.data_as(P_dbl) if a is not None else None


def _pl(a):
    return a.ctypes.data_as(P_ll) if a is not None else None


class Ref:
    """The reference itself (oracle/_ref/libvrref*.so)."""

    def __init__(self, variant: str = "serial"):
        name = {"serial": "libvrref.so", "omp": "libvrref_omp.so", "pp": "libvrref_pp.so", "gpu": "libvrref_gpu.so"}[variant]
        path = os.path.join(HERE, "_ref", name)
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (run `make -C oracle ref` where /root/reference exists)")
        self.variant = variant
        self.lib = C.CDLL(path)
        L = self.lib
        L.vrref_sizeof_particle.restype = C.c_size_t
        L.vrref_kdfof.restype = c_ll
        L.vrref_kdfof.argtypes = [c_ll, P_dbl, P_dbl, C.c_int, c_dbl, P_dbl, c_ll, C.c_int, P_ll]
        L.vrref_search_full_set.restype = c_ll
        L.vrref_search_full_set.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, P_ll, C.POINTER(RefOpts), P_ll, P_dbl]
        L.vrref_potential.restype = None
        L.vrref_potential.argtypes = [c_ll, P_dbl, P_dbl, C.POINTER(RefOpts), P_dbl]
        L.vrref_unbind.restype = c_ll
        L.vrref_unbind.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, C.POINTER(RefOpts), P_ll, c_ll,
                                   C.POINTER(C.c_int), P_dbl]
        if hasattr(L, "vrref_binding_energy"):
            L.vrref_binding_energy.restype = None
            L.vrref_binding_energy.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, c_ll, P_ll, P_dbl, C.POINTER(RefOpts), P_dbl, P_dbl,
                                               P_dbl, P_dbl, P_dbl, P_ll, P_ll]
        if hasattr(L, "vrref_sort_by_binding_energy"):
            L.vrref_sort_by_binding_energy.restype = None
            L.vrref_sort_by_binding_energy.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, P_ll, P_ll, c_ll, C.POINTER(RefOpts), P_ll, P_dbl,
                                                       P_ll, P_ll]
        if hasattr(L, "vrref_build_pglist"):
            L.vrref_build_pglist.restype = c_ll
            L.vrref_build_pglist.argtypes = [c_ll, P_ll, c_ll, P_ll, P_ll]
        assert L.vrref_sizeof_particle() == 96

    @staticmethod
    def available(variant: str = "serial") -> bool:
        name = {"serial": "libvrref.so", "omp": "libvrref_omp.so", "pp": "libvrref_pp.so", "gpu": "libvrref_gpu.so"}[variant]
        return os.path.exists(os.path.join(HERE, "_ref", name))

    def unbindnum(self) -> int:
        return int(self.lib.vrref_unbindnum())

    def kdfof(self, pos, vel=None, ndim=3, fdist=1.0, period=None, minnum=2, bucket=16):
        pos = _f64(pos)
        n = pos.shape[0]
        vel = _f64(vel) if vel is not None else None
        per = _f64([period] * 3) if period else None
        out = np.zeros(n, dtype=np.int64)
        ng = self.lib.vrref_kdfof(n, _pd(pos), _pd(vel), ndim, fdist, _pd(per), minnum, bucket, _pl(out))
        return out, int(ng)

    def fofcriterion6d(self, pos, vel, pot, p6, p7, p9=0.5, minsize=20, bucket=16):
        """KDTree::FOFCriterion(FOF6d, ..., FOFcheckbg) as called by the 6D core search (search.cxx:1607-1611)."""
        pos = _f64(pos); vel = _f64(vel); pot = _f64(pot)
        n = pos.shape[0]
        out = np.zeros(n, dtype=np.int64)
        f = self.lib.vrref_fofcriterion6d
        f.restype = c_ll
        f.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, c_dbl, c_dbl, c_dbl, c_ll, C.c_int, P_ll]
        ng = f(n, _pd(pos), _pd(vel), _pd(pot), p6, p7, p9, minsize, bucket, _pl(out))
        return out, int(ng)

    def search_full_set(self, pos, vel, mass, pid, opts: RefOpts):
        pos = _f64(pos).copy()
        vel = _f64(vel).copy()
        mass = _f64(mass)
        pid = _i64(pid)
        n = pos.shape[0]
        out = np.zeros(n, dtype=np.int64)
        sec = c_dbl(0)
        ng = self.lib.vrref_search_full_set(n, _pd(pos), _pd(vel), _pd(mass), _pl(pid), C.byref(opts), _pl(out),
                                            C.byref(sec))
        return {"pfof": out, "ngroups": int(ng), "pos": pos, "vel": vel, "seconds": sec.value}

    def velocity_density(self, pos, vel, period=0.0, bsize=16, nsearch=256, nvel=32, approx=1, ptype=None):
        """GetVelocityDensity (localfield.cxx:18) with tree == NULL; densities in input order."""
        pos = _f64(pos); vel = _f64(vel)
        n = pos.shape[0]
        den = np.zeros(n)
        tp = None if ptype is None else np.ascontiguousarray(ptype, dtype=np.int32)
        self.lib.vrref_velocity_density.restype = None
        self.lib.vrref_velocity_density(c_ll(n), _pd(pos), _pd(vel), None if tp is None else tp.ctypes.data_as(C.POINTER(C.c_int)),
                                        c_dbl(period), C.c_int(bsize), C.c_int(nsearch), C.c_int(nvel), C.c_int(approx), _pd(den))
        return den

    def potential(self, pos, mass, opts: RefOpts):
        pos = _f64(pos)
        mass = _f64(mass)
        n = pos.shape[0]
        phi = np.zeros(n)
        self.lib.vrref_potential(n, _pd(pos), _pd(mass), C.byref(opts), _pd(phi))
        return phi

    def unbind(self, pos, vel, mass, pfof, ngroup, opts: RefOpts):
        pos = _f64(pos)
        vel = _f64(vel)
        mass = _f64(mass)
        pf = _i64(pfof).copy()
        n = pos.shape[0]
        ch = C.c_int(0)
        sec = c_dbl(0)
        ng = self.lib.vrref_unbind(n, _pd(pos), _pd(vel), _pd(mass), C.byref(opts), _pl(pf), ngroup, C.byref(ch),
                                   C.byref(sec))
        return {"pfof": pf, "ngroups": int(ng), "changed": int(ch.value), "seconds": sec.value}


    def unbind_full(self, pos, vel, mass, pfof, ngroup, opts: RefOpts, ireorder=1, groupflag=None):
        """CheckUnboundGroups as SearchSubSub calls it (numingroup / pglist given, ireorder, optional groupflag)."""
        pos = _f64(pos); vel = _f64(vel); mass = _f64(mass)
        pf = _i64(pfof).copy()
        n = pos.shape[0]
        gf = None if groupflag is None else _i64(groupflag).copy()
        num = np.zeros(ngroup + 1, dtype=np.int64)
        ch = C.c_int(0)
        f = self.lib.vrref_unbind_full
        f.restype = c_ll
        ng = f(c_ll(n), _pd(pos), _pd(vel), _pd(mass), C.byref(opts), _pl(pf), c_ll(ngroup), C.c_int(ireorder),
               None if gf is None else _pl(gf), _pl(num), C.byref(ch))
        return {"pfof": pf, "ngroups": int(ng), "changed": int(ch.value), "numingroup": num, "groupflag": gf}

    def binding_energy(self, pos, vel, mass, numingroup, gcmvel, opts: RefOpts):
        """GetBindingEnergy on group-ordered particles; numingroup[g], gcmvel[g] for g = 1..G (index 0 unused)."""
        pos = _f64(pos); vel = _f64(vel); mass = _f64(mass)
        num = _i64(numingroup)
        gc = _f64(gcmvel)
        n, G = pos.shape[0], num.shape[0] - 1
        phi = np.zeros(n); E = np.zeros(n)
        T = np.zeros(G + 1); Pot = np.zeros(G + 1); Ef = np.zeros(G + 1)
        imin = np.zeros(G + 1, dtype=np.int64); ib = np.zeros(G + 1, dtype=np.int64)
        self.lib.vrref_binding_energy(n, _pd(pos), _pd(vel), _pd(mass), G, _pl(num), _pd(gc), C.byref(opts), _pd(phi), _pd(E),
                                      _pd(T), _pd(Pot), _pd(Ef), _pl(imin), _pl(ib))
        return {"phi": phi, "E": E, "T": T, "Pot": Pot, "Efrac": Ef, "iminpot": imin, "ibound": ib}


    def sort_by_binding_energy(self, pos, vel, mass, pid, pfof, ngroups, opts: RefOpts):
        """SortAccordingtoBindingEnergy (substructureproperties.cxx:4256): per group the PIDs / energies in the final particle
        order, numingroup[0..G] and iunbound[0..G] (pglist[g][numingroup[g]])."""
        pos = _f64(pos); vel = _f64(vel); mass = _f64(mass)
        pid = _i64(pid); pf = _i64(pfof)
        n = pos.shape[0]
        ps = np.zeros(n, dtype=np.int64); Es = np.zeros(n)
        num = np.zeros(ngroups + 1, dtype=np.int64); iu = np.zeros(ngroups + 1, dtype=np.int64)
        self.lib.vrref_sort_by_binding_energy(n, _pd(pos), _pd(vel), _pd(mass), _pl(pid), _pl(pf), ngroups, C.byref(opts), _pl(ps),
                                              _pd(Es), _pl(num), _pl(iu))
        ntot = int(num.sum())
        return {"pid": ps[:ntot], "E": Es[:ntot], "numingroup": num, "iunbound": iu}


    def build_pglist(self, pfof, ngroups):
        """BuildNumInGroup + BuildPGList (buildandsortarrays.cxx:11, 26): (numingroup[0..G], concatenated member lists)."""
        pf = _i64(pfof)
        num = np.zeros(ngroups + 1, dtype=np.int64)
        index = np.zeros(max(pf.shape[0], 1), dtype=np.int64)
        k = self.lib.vrref_build_pglist(pf.shape[0], _pl(pf), ngroups, _pl(num), _pl(index))
        return num, index[:k]


class UnbindOpts(C.Structure):
    """Mirror of vro_unbind_opts (oracle/vr_oracle.h)."""
    _fields_ = [
        ("G", c_dbl), ("eps", c_dbl), ("Eratio", c_dbl), ("minEfrac", c_dbl),
        ("unbindtype", C.c_int), ("bgpot", C.c_int),
        ("fracpotref", c_dbl), ("maxunbindfrac", c_dbl),
        ("maxunboundfracforiterativeunbind", c_dbl), ("maxallowedunboundfrac", c_dbl),
        ("theta", c_dbl), ("bucket", C.c_int),
        ("min_size", c_ll), ("pp_threshold", c_ll), ("large_threshold", c_ll),
        ("sortflag_mode", C.c_int), ("oldn_init", c_ll),
        ("cmvelreftype", C.c_int), ("npotref", c_ll),
    ]


def unbind_opts_from(o: RefOpts, pp_threshold: int = 150, sortflag_mode: int = 0, oldn_init: int = 0) -> UnbindOpts:
    u = UnbindOpts()
    u.G = o.G; u.eps = o.eps; u.Eratio = o.Eratio; u.minEfrac = o.minEfrac
    u.unbindtype = o.unbindtype; u.bgpot = o.bgpot; u.fracpotref = o.fracpotref
    u.maxunbindfrac = o.maxunbindfrac
    u.maxunboundfracforiterativeunbind = o.maxunboundfracforiterativeunbind
    u.maxallowedunboundfrac = o.maxallowedunboundfrac
    u.theta = o.TreeThetaOpen; u.bucket = o.BucketSize
    u.min_size = o.min_size; u.pp_threshold = pp_threshold; u.large_threshold = 1000
    u.sortflag_mode = sortflag_mode
    u.oldn_init = oldn_init
    u.cmvelreftype = o.cmvelreftype
    u.npotref = o.Npotref
    return u


class Oracle:
    """Our plain-C restatement (oracle/liboracle.so)."""

    def __init__(self):
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (run `make -C oracle oracle`)")
        self.lib = C.CDLL(path)
        L = self.lib
        L.vro_fof3d.restype = c_ll
        L.vro_fof3d.argtypes = [c_ll, P_dbl, c_dbl, P_dbl, c_ll, P_ll, P_ll]
        L.vro_adjust_for_period.restype = None
        L.vro_adjust_for_period.argtypes = [c_ll, P_dbl, c_dbl, P_ll, c_ll]
        L.vro_fof6d.restype = c_ll
        L.vro_fof6d.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, P_ll, c_ll, c_dbl, c_dbl, C.c_int, c_ll, P_ll, P_ll, P_dbl]
        L.vro_potential_pp.restype = None
        L.vro_potential_pp.argtypes = [c_ll, P_dbl, P_dbl, c_dbl, c_dbl, P_dbl]
        L.vro_potential_tree.restype = None
        L.vro_potential_tree.argtypes = [c_ll, P_dbl, P_dbl, c_dbl, c_dbl, c_dbl, C.c_int, P_dbl, P_ll]
        L.vro_unbind.restype = c_ll
        L.vro_unbind.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, P_ll, c_ll, C.POINTER(UnbindOpts), P_ll, c_dbl,
                                 P_ll, P_dbl, C.POINTER(C.c_int)]
        L.vro_binding_energy.restype = None
        L.vro_binding_energy.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, c_ll, P_ll, P_dbl, c_dbl, c_dbl, c_dbl, C.c_int, c_ll,
                                         P_dbl, P_dbl, P_dbl, P_dbl, P_dbl, P_ll, P_ll]

    def fof3d(self, pos, b, period=None, minsize=20, pid=None):
        pos = _f64(pos)
        n = pos.shape[0]
        per = _f64([period] * 3) if period else None
        pid = _i64(pid) if pid is not None else None
        out = np.zeros(n, dtype=np.int64)
        ng = self.lib.vro_fof3d(n, _pd(pos), b, _pd(per), minsize, _pl(pid), _pl(out))
        return out, int(ng)

    def fof_criterion6d(self, pos, vel, pot, p6, p7, p9=0.5, minsize=20, pid=None):
        pos = _f64(pos); vel = _f64(vel); pot = _f64(pot)
        n = pos.shape[0]
        pid = _i64(pid) if pid is not None else None
        out = np.zeros(n, dtype=np.int64)
        f = self.lib.vro_fof_criterion6d
        f.restype = c_ll
        f.argtypes = [c_ll, P_dbl, P_dbl, P_dbl, c_dbl, c_dbl, c_dbl, c_ll, P_ll, P_ll]
        ng = f(n, _pd(pos), _pd(vel), _pd(pot), p6, p7, p9, minsize, _pl(pid), _pl(out))
        return out, int(ng)

    def adjust_for_period(self, pos, period, pfof, ngroups):
        pos = _f64(pos).copy()
        pfof = _i64(pfof)
        self.lib.vro_adjust_for_period(pos.shape[0], _pd(pos), period, _pl(pfof), ngroups)
        return pos

    def fof6d(self, pos, vel, mass, pfof3d, ngroups3d, b6, vfac=1.25, mode=3, minsize=20, pid=None):
        pos = _f64(pos).copy(); vel = _f64(vel).copy(); mass = _f64(mass)
        pf3 = _i64(pfof3d)
        n = pos.shape[0]
        pid = _i64(pid) if pid is not None else None
        out = np.zeros(n, dtype=np.int64)
        vs2 = np.zeros(ngroups3d + 2)
        ng = self.lib.vro_fof6d(n, _pd(pos), _pd(vel), _pd(mass), _pl(pf3), ngroups3d, b6, vfac, mode, minsize,
                                _pl(pid), _pl(out), _pd(vs2))
        return {"pfof": out, "ngroups": int(ng), "pos": pos, "vel": vel, "vscale2": vs2[:ngroups3d + 1]}

    def potential_pp(self, pos, mass, G, eps=0.0):
        pos = _f64(pos); mass = _f64(mass)
        phi = np.zeros(pos.shape[0])
        self.lib.vro_potential_pp(pos.shape[0], _pd(pos), _pd(mass), G, eps, _pd(phi))
        return phi

    def velocity_density(self, pos, vel, period=0.0, bsize=16, nsearch=256, nvel=32, approx=1, ptype=None):
        """Row f4, oracle side: GetVelocityDensity for one structure; densities in input order."""
        pos = _f64(pos); vel = _f64(vel)
        n = pos.shape[0]
        den = np.zeros(n)
        tp = None if ptype is None else np.ascontiguousarray(ptype, dtype=np.int32)
        self.lib.vro_velocity_density.restype = c_ll
        rc = self.lib.vro_velocity_density(c_ll(n), _pd(pos), _pd(vel), None if tp is None else tp.ctypes.data_as(C.POINTER(C.c_int)),
                                           c_dbl(period), C.c_int(bsize), C.c_int(nsearch), C.c_int(nvel), C.c_int(approx), _pd(den))
        if rc < 0:
            raise ValueError("nsearch / nvel do not fit the particle count")
        return den

    def tree_order(self, pos, bucket=8):
        """(perm, nodes): tree position -> input index, and [start, end) of every node in pre-order."""
        pos = _f64(pos)
        n = pos.shape[0]
        perm = np.zeros(n, dtype=np.int64)
        cap = 4 * n // max(1, bucket) + 64
        nodes = np.zeros((cap, 2), dtype=np.int64)
        self.lib.vro_tree_order.restype = c_ll
        nn = self.lib.vro_tree_order(c_ll(n), _pd(pos), C.c_int(bucket), _pl(perm), _pl(nodes), c_ll(cap))
        return perm, nodes[:int(nn)]

    def potential_tree(self, pos, mass, G, eps=0.0, theta=0.5, bucket=8):
        pos = _f64(pos); mass = _f64(mass)
        phi = np.zeros(pos.shape[0])
        inter = c_ll(0)
        self.lib.vro_potential_tree(pos.shape[0], _pd(pos), _pd(mass), G, eps, theta, bucket, _pd(phi), C.byref(inter))
        return phi, int(inter.value)

    def unbind(self, pos, vel, mass, pfof, ngroups, uopts: UnbindOpts, pid=None, marginal_tol=1e-6):
        pos = _f64(pos); vel = _f64(vel); mass = _f64(mass)
        pf = _i64(pfof).copy()
        n = pos.shape[0]
        pid = _i64(pid) if pid is not None else None
        nm = c_ll(0)
        ch = C.c_int(0)
        phi = np.zeros(n)
        ng = self.lib.vro_unbind(n, _pd(pos), _pd(vel), _pd(mass), _pl(pf), ngroups, C.byref(uopts), _pl(pid),
                                 marginal_tol, C.byref(nm), _pd(phi), C.byref(ch))
        return {"pfof": pf, "ngroups": int(ng), "n_marginal": int(nm.value), "phi": phi, "changed": int(ch.value)}

    def binding_energy(self, pos, vel, mass, numingroup, gcmvel, G, eps=0.0, theta=0.5, bucket=8, large_threshold=1000):
        """GetBindingEnergy (substructureproperties.cxx:3884) on group-ordered particles (see vr_oracle.h)."""
        pos = _f64(pos); vel = _f64(vel); mass = _f64(mass)
        num = _i64(numingroup)
        gc = _f64(gcmvel)
        n, ng = pos.shape[0], num.shape[0] - 1
        off = np.zeros(ng + 2, dtype=np.int64)
        off[2:] = np.cumsum(num[1:])
        phi = np.zeros(n); E = np.zeros(n)
        T = np.zeros(ng + 1); Pot = np.zeros(ng + 1); Ef = np.zeros(ng + 1)
        imin = np.zeros(ng + 1, dtype=np.int64); ib = np.zeros(ng + 1, dtype=np.int64)
        self.lib.vro_binding_energy(n, _pd(pos), _pd(vel), _pd(mass), ng, _pl(off), _pd(gc), G, eps, theta, bucket,
                                    large_threshold, _pd(phi), _pd(E), _pd(T), _pd(Pot), _pd(Ef), _pl(imin), _pl(ib))
        return {"phi": phi, "E": E, "T": T, "Pot": Pot, "Efrac": Ef, "iminpot": imin, "ibound": ib}
