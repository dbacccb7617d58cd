"""ctypes binding of libvrgpu.so (include/vrgpu.h) -- the B200 structure-finding hot path.

This module is host-side plumbing only: it builds `vrgpu_view`s over numpy / torch host buffers and forwards
to the C ABI.  There is NO CPU fallback: if the CUDA library is missing or no B200 is present, loading or
`Context()` raises.

`PARTICLE_DTYPE` is the reference's default 96-byte NBody::Particle layout
(NBodylib/src/NBody/Particle.h:75-109: mass, position[3], velocity[3], pid, id, type, rho, phi).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libvrgpu.so")

PARTICLE_DTYPE = np.dtype({
    "names": ["mass", "pos", "vel", "pid", "id", "type", "rho", "phi"],
    "formats": ["<f8", ("<f8", 3), ("<f8", 3), "<i8", "<i8", "<i4", "<f8", "<f8"],
    "offsets": [0, 8, 32, 56, 64, 72, 80, 88],
    "itemsize": 96,
})

NOFIELD = C.c_size_t(-1).value

EINVAL = 3
NOALLOC_MAX_PARTICLES = 65536   # VRGPU_NOALLOC_MAX_PARTICLES (include/vrgpu.h)
ERRORS = {1: "VRGPU_ENODEV", 2: "VRGPU_ECUDA", 3: "VRGPU_EINVAL", 4: "VRGPU_EUNSUPPORTED", 5: "VRGPU_ENOMEM"}


class VrgpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{ERRORS.get(code, code)}: {msg}")
        self.code = code


class View(C.Structure):
    _fields_ = [("base", C.c_void_p), ("stride", C.c_size_t), ("off_pos", C.c_size_t), ("off_vel", C.c_size_t),
                ("off_mass", C.c_size_t), ("off_pid", C.c_size_t), ("n", C.c_longlong)]


class UnbindOpts(C.Structure):
    _fields_ = [("G", C.c_double), ("eps", C.c_double), ("Eratio", C.c_double), ("minEfrac", C.c_double),
                ("unbindtype", C.c_int), ("bgpot", C.c_int), ("fracpotref", C.c_double),
                ("maxunbindfrac", C.c_double), ("maxunboundfracforiterativeunbind", C.c_double),
                ("maxallowedunboundfrac", C.c_double), ("theta", C.c_double), ("bucket", C.c_int),
                ("min_size", C.c_longlong), ("pp_threshold", C.c_longlong), ("large_threshold", C.c_longlong),
                ("marginal_tol", C.c_double), ("cmvelreftype", C.c_int), ("npotref", C.c_longlong)]


class SearchOpts(C.Structure):
    _fields_ = [("period", C.c_double), ("param1", C.c_double), ("halo_min_size", C.c_longlong),
                ("fofbgtype", C.c_int), ("ellhalo6dxfac", C.c_double), ("ellhalo6dvfac", C.c_double),
                ("bound_halos", C.c_int), ("unbind", UnbindOpts)]


class Stats(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("h2d_ms", "sort_ms", "link_ms", "label_ms", "fof6d_ms", "potential_ms",
                                          "unbind_ms", "d2h_ms", "total_ms")] + \
               [(k, C.c_longlong) for k in ("n_cells", "n_groups3d", "n_groups_final", "n_grouped",
                                            "n_pairs_potential", "n_marginal", "h2d_bytes", "d2h_bytes",
                                            "kernel_launches", "n_tree_interactions", "n_tree_nodes", "n_link_queue")]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def default_unbind_opts(**kw) -> UnbindOpts:
    """UnbindInfo defaults (allvars.h:312-329) overlaid with the sample cfg's unbinding block
    (examples/sample_dmcosmological_run.cfg: Unbinding_type=0, Eratio 0.95, Min_bound_mass_frac 0.65)."""
    u = UnbindOpts()
    u.G = 43.0211349
    u.eps = 0.0
    u.Eratio = 0.95
    u.minEfrac = 0.65
    u.unbindtype = 0
    u.bgpot = 1
    u.fracpotref = 1.0
    u.maxunbindfrac = 0.5
    u.maxunboundfracforiterativeunbind = 0.95
    u.maxallowedunboundfrac = 0.025
    u.theta = 0.5
    u.bucket = 8
    u.min_size = 20
    u.pp_threshold = 150
    u.large_threshold = 1000
    u.marginal_tol = 1e-6
    u.cmvelreftype = 0
    u.npotref = 20
    for k, v in kw.items():
        if not hasattr(u, k):
            raise AttributeError(k)
        setattr(u, k, v)
    return u


_lib = None


def load_library(path: str = LIB_PATH):
    """Load libvrgpu.so; raises (never falls back) if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
    L = C.CDLL(path)
    L.vrgpu_version.restype = C.c_char_p
    L.vrgpu_init.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    L.vrgpu_destroy.argtypes = [C.c_void_p]
    L.vrgpu_destroy.restype = None
    L.vrgpu_last_error.argtypes = [C.c_void_p]
    L.vrgpu_last_error.restype = C.c_char_p
    L.vrgpu_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
    L.vrgpu_get_stats.restype = None
    P_ll = C.POINTER(C.c_longlong)
    P_d = C.POINTER(C.c_double)
    L.vrgpu_fof3d.argtypes = [C.c_void_p, View, C.c_double, P_d, C.c_longlong, P_ll, P_ll]
    L.vrgpu_adjust_for_period.argtypes = [C.c_void_p, View, C.c_double, P_ll, C.c_longlong]
    L.vrgpu_fof6d.argtypes = [C.c_void_p, View, P_ll, C.c_longlong, C.c_double, C.c_double, C.c_int, C.c_longlong, P_ll, P_ll]
    L.vrgpu_potential.argtypes = [C.c_void_p, View, C.c_longlong, P_ll, P_ll, C.c_double, C.c_double, C.c_double, C.c_int,
                                  C.c_longlong, P_d]
    L.vrgpu_tree_order.argtypes = [C.c_void_p, View, C.c_longlong, P_ll, P_ll, C.c_int, P_ll]
    L.vrgpu_binding_energy.argtypes = [C.c_void_p, View, C.c_longlong, P_ll, P_d, C.c_double, C.c_double, C.c_double, C.c_int,
                                       C.c_longlong, P_d, P_d, P_d, P_d, P_d, P_ll, P_ll]
    L.vrgpu_sort_groups_by_key.argtypes = [C.c_void_p, C.c_longlong, P_ll, P_d, P_ll]
    L.vrgpu_build_pglist.argtypes = [C.c_void_p, P_ll, C.c_longlong, C.c_longlong, P_ll, P_ll, P_ll]
    L.vrgpu_unbind.argtypes = [C.c_void_p, View, P_ll, P_ll, C.POINTER(UnbindOpts), C.POINTER(C.c_int), P_ll]
    L.vrgpu_search_full_set.argtypes = [C.c_void_p, View, C.POINTER(SearchOpts), P_ll, P_ll]
    L.vrgpu_particles_upload.argtypes = [C.c_void_p, View]
    L.vrgpu_particles_release.argtypes = [C.c_void_p]
    L.vrgpu_search_full_set_resident.argtypes = [C.c_void_p, C.POINTER(SearchOpts), P_ll]
    L.vrgpu_resident_pfof.argtypes = [C.c_void_p, P_ll]
    L.vrgpu_dist_unique_id.argtypes = [C.c_char_p]
    L.vrgpu_dist_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
    L.vrgpu_dist_search_full_set.argtypes = [C.c_void_p, View, C.c_double, C.c_double, C.POINTER(SearchOpts), P_ll, P_ll]
    L.vrgpu_dist_search_full_set_resident.argtypes = [C.c_void_p, C.c_double, C.c_double, C.POINTER(SearchOpts), P_ll]
    L.vrgpu_fof_criterion6d.argtypes = [C.c_void_p, View, P_d, C.c_double, C.c_double, C.c_double, C.c_longlong, P_ll, P_ll]
    L.vrgpu_velocity_density.argtypes = [C.c_void_p, View, C.POINTER(C.c_int), P_ll, C.c_longlong, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, P_d]
    L.vrgpu_stream.argtypes = [C.c_void_p]
    L.vrgpu_stream.restype = C.c_void_p
    L.vrgpu_kernel_times.argtypes = [C.c_void_p]
    L.vrgpu_kernel_times.restype = C.c_char_p
    L.vrgpu_measure_fp64_tflops.argtypes = [C.c_void_p, P_d]
    L.vrgpu_probe_fp64_pipes.argtypes = [C.c_void_p, P_d]
    L.vrgpu_kernel_times_reset.argtypes = [C.c_void_p]
    # device pointers as integers (c_void_p): pos_dev, mass_dev, phi_dev
    L.vrgpu_potential_direct_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_double, C.c_double,
                                                 C.c_void_p, C.c_char_p, C.c_longlong]
    L.vrgpu_potential_direct_noalloc.restype = C.c_int
    L.vrgpu_kernel_times_reset.restype = None
    _lib = L
    return L


EXPORTED_SYMBOLS = ["vrgpu_init", "vrgpu_destroy", "vrgpu_last_error", "vrgpu_version", "vrgpu_get_stats", "vrgpu_fof3d",
                    "vrgpu_adjust_for_period", "vrgpu_fof6d", "vrgpu_potential", "vrgpu_unbind", "vrgpu_search_full_set",
                    "vrgpu_particles_upload", "vrgpu_particles_release", "vrgpu_search_full_set_resident",
                    "vrgpu_resident_pfof", "vrgpu_dist_unique_id", "vrgpu_dist_init", "vrgpu_dist_search_full_set",
                    "vrgpu_dist_search_full_set_resident", "vrgpu_stream", "vrgpu_kernel_times", "vrgpu_kernel_times_reset", "vrgpu_measure_fp64_tflops", "vrgpu_probe_fp64_pipes", "vrgpu_fof_criterion6d", "vrgpu_velocity_density",
                    "vrgpu_binding_energy", "vrgpu_sort_groups_by_key", "vrgpu_build_pglist", "vrgpu_tree_order", "vrgpu_potential_direct_noalloc"]


def pinned_i64(n: int) -> np.ndarray:
    """A page-locked int64 host array (through torch), e.g. a reusable pfof_out buffer for the search calls."""
    import torch
    t = torch.empty(max(int(n), 1), dtype=torch.int64).pin_memory()
    a = t.numpy()[:n]
    _pinned_keepalive.append(t)
    return a


_pinned_keepalive = []


def particles_from_arrays(pos, vel=None, mass=None, pid=None, pinned: bool = False) -> np.ndarray:
    """Build the reference's 96-byte AoS (what SearchFullSet receives as vector<Particle>) from columns.
    pinned=True allocates the records in page-locked host memory through torch (zero-copy upload path)."""
    pos = np.asarray(pos, dtype=np.float64)
    n = pos.shape[0]
    if pinned:
        import torch
        buf = torch.empty(n * PARTICLE_DTYPE.itemsize, dtype=torch.uint8, pin_memory=True)
        parts = buf.numpy().view(PARTICLE_DTYPE)
        _KEEP.append(buf)
    else:
        parts = np.zeros(n, dtype=PARTICLE_DTYPE)
    parts["pos"] = pos
    parts["vel"] = 0.0 if vel is None else np.asarray(vel, dtype=np.float64)
    parts["mass"] = 1.0 if mass is None else np.asarray(mass, dtype=np.float64)
    parts["pid"] = np.arange(n, dtype=np.int64) if pid is None else np.asarray(pid, dtype=np.int64)
    parts["id"] = np.arange(n, dtype=np.int64)
    parts["type"] = 1
    parts["rho"] = 0.0
    parts["phi"] = 0.0
    return parts


_KEEP: list = []


def view_of(parts: np.ndarray) -> View:
    """vrgpu_view over a PARTICLE_DTYPE array (no copy)."""
    assert parts.dtype == PARTICLE_DTYPE and parts.flags["C_CONTIGUOUS"]
    v = View()
    v.base = parts.ctypes.data
    v.stride = PARTICLE_DTYPE.itemsize
    v.off_pos, v.off_vel, v.off_mass, v.off_pid = 8, 32, 0, 56
    v.n = parts.shape[0]
    return v


class Context:
    """One vrgpu_ctx (one CUDA device, one caller thread)."""

    def __init__(self, device: int = -1):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.vrgpu_init(device, C.byref(h))
        if rc != 0:
            raise VrgpuError(rc, "vrgpu_init failed (no B200 / CUDA device available; there is no CPU fallback)")
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.vrgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise VrgpuError(rc, self.lib.vrgpu_last_error(self.h).decode())

    def stats(self) -> dict:
        s = Stats()
        self.lib.vrgpu_get_stats(self.h, C.byref(s))
        return s.asdict()

    # ---- stage calls (names follow the reference functions they replace) ----
    def fof3d(self, parts, b, period=None, minsize=20, out=None):
        """KDTree::FOF (KDFOF.cxx:12) as used by SearchFullSet (search.cxx:203)."""
        n = parts.shape[0]
        pfof = out if out is not None else np.zeros(n, dtype=np.int64)
        ng = C.c_longlong(0)
        per = (C.c_double * 3)(period, period, period) if period else None
        self._check(self.lib.vrgpu_fof3d(self.h, view_of(parts), b, per, minsize,
                                         pfof.ctypes.data_as(C.POINTER(C.c_longlong)), C.byref(ng)))
        return pfof, int(ng.value)

    def adjust_for_period(self, parts, period, pfof, ngroups):
        """AdjustStructureForPeriod (search.cxx:856); parts['pos'] is updated in place."""
        pfof = np.ascontiguousarray(pfof, dtype=np.int64)
        self._check(self.lib.vrgpu_adjust_for_period(self.h, view_of(parts), period,
                                                     pfof.ctypes.data_as(C.POINTER(C.c_longlong)), ngroups))

    def fof6d(self, parts, pfof3d, ngroups3d, b6, vfac=1.25, mode=3, minsize=20):
        """6DFOF block of SearchFullSet (search.cxx:400-698); parts pos/vel get the Q3 rounding in place."""
        n = parts.shape[0]
        pfof3d = np.ascontiguousarray(pfof3d, dtype=np.int64)
        pfof = np.zeros(n, dtype=np.int64)
        ng = C.c_longlong(0)
        P = C.POINTER(C.c_longlong)
        self._check(self.lib.vrgpu_fof6d(self.h, view_of(parts), pfof3d.ctypes.data_as(P), ngroups3d, b6, vfac, mode,
                                         minsize, pfof.ctypes.data_as(P), C.byref(ng)))
        return pfof, int(ng.value)

    def potential(self, parts, offsets, index, G, eps=0.0, theta=0.5, bucket=8, pp_threshold=150):
        """Unbind's direct sum (unbind.cxx:398-416) / Potential() (unbind.cxx:1607) for CSR groups."""
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        index = np.ascontiguousarray(index, dtype=np.int64)
        ng = offsets.shape[0] - 1
        phi = np.zeros(index.shape[0], dtype=np.float64)
        P = C.POINTER(C.c_longlong)
        self._check(self.lib.vrgpu_potential(self.h, view_of(parts), ng, offsets.ctypes.data_as(P), index.ctypes.data_as(P),
                                             G, eps, theta, bucket, pp_threshold, phi.ctypes.data_as(C.POINTER(C.c_double))))
        return phi

    def tree_order(self, parts, offsets, index, bucket=8):
        """The arrangement KDTree::BuildNodes / MedianPos (KDTree.cxx:438-507, 304-337) leave in every group: perm[offsets[g] + j] =
        offset inside group g of the member at tree position j."""
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        index = np.ascontiguousarray(index, dtype=np.int64)
        perm = np.zeros(index.shape[0], dtype=np.int64)
        P = C.POINTER(C.c_longlong)
        self._check(self.lib.vrgpu_tree_order(self.h, view_of(parts), offsets.shape[0] - 1, offsets.ctypes.data_as(P), index.ctypes.data_as(P),
                                              bucket, perm.ctypes.data_as(P)))
        return perm

    def binding_energy(self, parts, offsets, gcmvel, G, eps=0.0, theta=0.5, bucket=8, large_threshold=1000):
        """GetBindingEnergy (substructureproperties.cxx:3884-4255) for group-ordered particles; group g = parts[offsets[g]:offsets[g+1]]."""
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        gcmvel = np.ascontiguousarray(gcmvel, dtype=np.float64)
        ng = offsets.shape[0] - 1
        ntot = int(offsets[-1])
        r = {"phi": np.zeros(ntot), "E": np.zeros(ntot), "T": np.zeros(ng), "Pot": np.zeros(ng), "Efrac": np.zeros(ng),
             "iminpot": np.zeros(ng, dtype=np.int64), "ibound": np.zeros(ng, dtype=np.int64)}
        D, LL = C.POINTER(C.c_double), C.POINTER(C.c_longlong)
        self._check(self.lib.vrgpu_binding_energy(self.h, view_of(parts), ng, offsets.ctypes.data_as(LL), gcmvel.ctypes.data_as(D), G, eps,
                                                  theta, bucket, large_threshold, r["phi"].ctypes.data_as(D), r["E"].ctypes.data_as(D),
                                                  r["T"].ctypes.data_as(D), r["Pot"].ctypes.data_as(D), r["Efrac"].ctypes.data_as(D),
                                                  r["iminpot"].ctypes.data_as(LL), r["ibound"].ctypes.data_as(LL)))
        return r

    def sort_groups_by_key(self, offsets, key):
        """Per-group ascending order of key (the qsort of SortAccordingtoBindingEnergy, substructureproperties.cxx:4302-4310)."""
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        key = np.ascontiguousarray(key, dtype=np.float64)
        order = np.zeros(int(offsets[-1]), dtype=np.int64)
        self._check(self.lib.vrgpu_sort_groups_by_key(self.h, offsets.shape[0] - 1, offsets.ctypes.data_as(C.POINTER(C.c_longlong)),
                                                      key.ctypes.data_as(C.POINTER(C.c_double)), order.ctypes.data_as(C.POINTER(C.c_longlong))))
        return order

    def build_pglist(self, pfof=None, ngroups=0, n_resident=0):
        """BuildNumInGroup + BuildPGList (buildandsortarrays.cxx:11-16, 26-43): returns (numingroup[0..G], index) with the
        members of group 1, 2, ... concatenated in ascending particle index.  pfof=None: the resident labels (pass the
        resident particle count as n_resident and the group count as ngroups to size the outputs)."""
        LL = C.POINTER(C.c_longlong)
        if pfof is None:
            n, pf = int(n_resident), None
        else:
            pf = np.ascontiguousarray(pfof, dtype=np.int64)
            n = pf.shape[0]
        num = np.zeros(int(ngroups) + 1, dtype=np.int64)
        index = np.zeros(max(n, 1), dtype=np.int64)
        ntot = C.c_longlong(0)
        self._check(self.lib.vrgpu_build_pglist(self.h, pf.ctypes.data_as(LL) if pf is not None else None, n, int(ngroups),
                                                num.ctypes.data_as(LL), index.ctypes.data_as(LL), C.byref(ntot)))
        return num, index[:int(ntot.value)]

    def fof_criterion6d(self, parts, check, xlink2, vlink2, check_below=0.5, minsize=20):
        """KDTree::FOFCriterion(FOF6d, ..., FOFcheckbg): the linking step of the 6D core search (search.cxx:1607-1611)."""
        n = parts.shape[0]
        pf = np.zeros(n, dtype=np.int64)
        ng = C.c_longlong(0)
        chk = None if check is None else np.ascontiguousarray(check, dtype=np.float64)
        self._check(self.lib.vrgpu_fof_criterion6d(self.h, view_of(parts), None if chk is None else chk.ctypes.data_as(C.POINTER(C.c_double)),
                                                   check_below, xlink2, vlink2, minsize, pf.ctypes.data_as(C.POINTER(C.c_longlong)), C.byref(ng)))
        return pf, int(ng.value)

    def velocity_density(self, parts, period=0.0, bsize=16, nsearch=256, nvel=32, approx=1, ptype=None, leaf_offsets=None):
        """GetVelocityDensity (localfield.cxx:18) for the particles of one structure; densities by particle index.
        leaf_offsets: leaves of the caller's own tree (particles in that tree's order), see include/vrgpu.h."""
        n = parts.shape[0]
        den = np.zeros(n, dtype=np.float64)
        tp = None if ptype is None else np.ascontiguousarray(ptype, dtype=np.int32)
        lo = None if leaf_offsets is None else np.ascontiguousarray(leaf_offsets, dtype=np.int64)
        self._check(self.lib.vrgpu_velocity_density(self.h, view_of(parts), None if tp is None else tp.ctypes.data_as(C.POINTER(C.c_int)),
                                                    None if lo is None else lo.ctypes.data_as(C.POINTER(C.c_longlong)),
                                                    0 if lo is None else len(lo) - 1, float(period or 0.0), bsize, nsearch, nvel, approx,
                                                    den.ctypes.data_as(C.POINTER(C.c_double))))
        return den

    def unbind(self, parts, pfof, ngroups, opts: UnbindOpts):
        """CheckUnboundGroups (unbind.cxx:196)."""
        pf = np.ascontiguousarray(pfof, dtype=np.int64).copy()
        ng = C.c_longlong(ngroups)
        ch = C.c_int(0)
        nm = C.c_longlong(0)
        self._check(self.lib.vrgpu_unbind(self.h, view_of(parts), pf.ctypes.data_as(C.POINTER(C.c_longlong)), C.byref(ng),
                                          C.byref(opts), C.byref(ch), C.byref(nm)))
        return {"pfof": pf, "ngroups": int(ng.value), "changed": int(ch.value), "n_marginal": int(nm.value)}

    def search_full_set(self, parts, opts: SearchOpts, out=None):
        """SearchFullSet (search.cxx:24-854), DM-only path."""
        n = parts.shape[0]
        pfof = out if out is not None else np.zeros(n, dtype=np.int64)
        ng = C.c_longlong(0)
        self._check(self.lib.vrgpu_search_full_set(self.h, view_of(parts), C.byref(opts),
                                                   pfof.ctypes.data_as(C.POINTER(C.c_longlong)), C.byref(ng)))
        return pfof, int(ng.value)

    # ---- device-resident variant (InvokeVelociraptor-style: particles handed over once) ----
    def particles_upload(self, parts):
        self._check(self.lib.vrgpu_particles_upload(self.h, view_of(parts)))

    def particles_release(self):
        self._check(self.lib.vrgpu_particles_release(self.h))

    def search_full_set_resident(self, opts: SearchOpts) -> int:
        ng = C.c_longlong(0)
        self._check(self.lib.vrgpu_search_full_set_resident(self.h, C.byref(opts), C.byref(ng)))
        return int(ng.value)

    def resident_pfof(self, n: int) -> np.ndarray:
        pfof = np.zeros(n, dtype=np.int64)
        self._check(self.lib.vrgpu_resident_pfof(self.h, pfof.ctypes.data_as(C.POINTER(C.c_longlong))))
        return pfof

    # ---- multi-GPU (one process per GPU; replaces the MPI build of SearchFullSet) ----
    def dist_init(self, nranks: int, rank: int, unique_id: bytes | None):
        self._check(self.lib.vrgpu_dist_init(self.h, nranks, rank, unique_id))

    def dist_search_full_set(self, parts, slab_lo, slab_hi, opts: SearchOpts, out=None):
        n = parts.shape[0]
        pfof = out if out is not None else np.zeros(n, dtype=np.int64)
        ng = C.c_longlong(0)
        self._check(self.lib.vrgpu_dist_search_full_set(self.h, view_of(parts), slab_lo, slab_hi, C.byref(opts),
                                                        pfof.ctypes.data_as(C.POINTER(C.c_longlong)), C.byref(ng)))
        return pfof, int(ng.value)

    def dist_search_full_set_resident(self, slab_lo, slab_hi, opts: SearchOpts) -> int:
        ng = C.c_longlong(0)
        self._check(self.lib.vrgpu_dist_search_full_set_resident(self.h, slab_lo, slab_hi, C.byref(opts), C.byref(ng)))
        return int(ng.value)

    # ---- measurement hooks ----
    def stream(self) -> int:
        return int(self.lib.vrgpu_stream(self.h) or 0)

    def kernel_times(self) -> dict:
        out = {}
        for line in self.lib.vrgpu_kernel_times(self.h).decode().splitlines():
            name, ms, cnt = line.split()
            out[name] = {"ms": float(ms), "count": int(cnt)}
        return out

    def measure_fp64_tflops(self) -> float:
        v = C.c_double(0)
        self._check(self.lib.vrgpu_measure_fp64_tflops(self.h, C.byref(v)))
        return float(v.value)

    def probe_fp64_pipes(self) -> dict:
        v = (C.c_double * 4)()
        self._check(self.lib.vrgpu_probe_fp64_pipes(self.h, v))
        return {"dfma_alone": v[0], "dmma_alone": v[1], "dfma_mixed": v[2], "dmma_mixed": v[3]}

    def kernel_times_reset(self):
        self.lib.vrgpu_kernel_times_reset(self.h)


def dist_unique_id() -> bytes:
    """128-byte NCCL id created on rank 0; broadcast it to every rank before Context.dist_init."""
    buf = C.create_string_buffer(128)
    rc = load_library().vrgpu_dist_unique_id(buf)
    if rc != 0:
        raise VrgpuError(rc, "vrgpu_dist_unique_id failed (NCCL unavailable?)")
    return buf.raw


def search_opts(period, param1, halo_min_size=20, fofbgtype=5, ellhalo6dxfac=1.0, ellhalo6dvfac=1.25, bound_halos=0,
                unbind: UnbindOpts | None = None) -> SearchOpts:
    o = SearchOpts()
    o.period = period
    o.param1 = param1
    o.halo_min_size = halo_min_size
    o.fofbgtype = fofbgtype
    o.ellhalo6dxfac = ellhalo6dxfac
    o.ellhalo6dvfac = ellhalo6dvfac
    o.bound_halos = bound_halos
    o.unbind = unbind if unbind is not None else default_unbind_opts()
    return o
