/* include/vrgpu.h -- C ABI of libvrgpu.so, the B200 (sm_100a) implementation of VELOCIraptor's
 * structure-finding hot path: 3D friends-of-friends -> 6D phase-space FOF -> potential + unbinding.
 *
 * This is the drop-in boundary.  Every entry point replaces one reference interface (paths under the
 * reference tree, cited per function) and is what the reference-side host adaptor
 * (velociraptor-stf_b200/host/vrgpu_adaptor.cxx) binds.
 *
 * Conventions
 *   - plain C, no exceptions cross the boundary; return 0 = OK, non-zero = VRGPU_E* code,
 *     vrgpu_last_error(ctx) gives the text.
 *   - every pointer is a HOST pointer owned by the caller (pinned or pageable); the library never frees
 *     or retains caller memory (one exception, stated there: vrgpu_potential_direct_noalloc takes device pointers).  Outputs are fully written on success and untouched on failure.
 *   - Int_t == long long (the reference's default LONGINT build, NBodylib/src/Math/Precision.h:49),
 *     Double_t == double.  NBody::Particle must be the 96-byte default layout
 *     (NBodylib/src/NBody/Particle.h:75-109); other layouts are described through vrgpu_view.
 *   - one caller thread per context (all reference call sites are on the main thread).
 *   - there is NO CPU fallback: without a CUDA device vrgpu_init fails with VRGPU_ENODEV.
 */
#ifndef VRGPU_H
#define VRGPU_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VRGPU_OK 0
#define VRGPU_ENODEV 1      /* no usable CUDA device */
#define VRGPU_ECUDA 2       /* a CUDA runtime call failed */
#define VRGPU_EINVAL 3      /* bad argument */
#define VRGPU_EUNSUPPORTED 4/* option combination outside the implemented path */
#define VRGPU_ENOMEM 5

typedef struct vrgpu_ctx vrgpu_ctx;

/* Strided view straight into the caller's particle array (AoS NBody::Particle or anything else), so the
 * host does no repacking.  Offsets are in bytes from `base`; pos/vel are 3 consecutive doubles, mass one
 * double, pid one long long.  off_vel/off_mass/off_pid may be (size_t)-1 when a call does not need them
 * (mass then defaults to 1.0, pid to the index). */
typedef struct {
    const void *base;
    size_t stride;
    size_t off_pos, off_vel, off_mass, off_pid;
    long long n;
} vrgpu_view;
#define VRGPU_NOFIELD ((size_t)-1)

/* Unbinding controls == reference UnbindInfo (src/allvars.h:270-330) + the Options members Unbind() reads. */
typedef struct {
    double G;                  /* Options::G                                                        */
    double eps;                /* UnbindInfo::eps (softening, code length)                          */
    double Eratio;             /* UnbindInfo::Eratio                                                */
    double minEfrac;           /* UnbindInfo::minEfrac                                              */
    int unbindtype;            /* UnbindInfo::unbindtype 0 = USYSANDPART, 1 = UPART                 */
    int bgpot;                 /* UnbindInfo::bgpot; 1 keep the background potential, 0 take removed particles' pair terms off */
    double fracpotref;         /* UnbindInfo::fracpotref: 1.0 = CM velocity of the whole group, else the frame of the innermost
                                  max(npotref, fracpotref * n) members (unbind.cxx:460-544), see cmvelreftype */
    double maxunbindfrac;      /* UnbindInfo::maxunbindfrac (0.5)                                   */
    double maxunboundfracforiterativeunbind; /* (0.95)                                              */
    double maxallowedunboundfrac;            /* (0.025)                                             */
    double theta;              /* UnbindInfo::TreeThetaOpen (0.5)                                   */
    int bucket;                /* UnbindInfo::BucketSize (8)                                        */
    long long min_size;        /* Options::MinSize at the call                                      */
    long long pp_threshold;    /* UNBINDNUM (150): n <= this -> direct sum; < 0 -> direct sum always */
    long long large_threshold; /* ompunbindnum (1000): selects the reference's two loop variants    */
    double marginal_tol;       /* |E| <= tol*|phi| counts as marginal (BASELINE.json: 1e-6)         */
    int cmvelreftype;          /* UnbindInfo::cmvelreftype, read when fracpotref != 1: 0 CMVELREF (radii about the centre of mass),
                                  1 POTREF (about the member with the lowest potential) -- allvars.h:216-218 */
    long long npotref;         /* UnbindInfo::Npotref (20)                                          */
} vrgpu_unbind_opts;

/* Options subset SearchFullSet reads on the DM-only path (src/search.cxx:24-854). */
typedef struct {
    double period;             /* Options::p, 0 = open                                              */
    double param1;             /* param[1] = (ellxscale^2)*(ellphys^2)*(ellhalophysfac^2), the SQUARED physical linking
                                  length exactly as search.cxx:77 forms it; 3DFOF uses b = sqrt(param1), b2 = b*b
                                  (search.cxx:203, KDFOF.cxx:14); 6DFOF uses param1*ellhalo6dxfac^2 (search.cxx:410) */
    long long halo_min_size;   /* Options::HaloMinSize                                              */
    int fofbgtype;             /* 5 FOF3D, 3 FOF6DADAPTIVE, 4 FOF6D (Options::fofbgtype)            */
    double ellhalo6dxfac;      /* Options::ellhalo6dxfac (search.cxx:410)                           */
    double ellhalo6dvfac;      /* Options::ellhalo6dvfac (search.cxx:464,492)                       */
    int bound_halos;           /* Options::iBoundHalos >= 1 -> CheckUnboundGroups (search.cxx:702)  */
    vrgpu_unbind_opts unbind;
} vrgpu_search_opts;

/* Per-stage device timings (milliseconds, CUDA events) of the last call, for bench.py / profiling. */
typedef struct {
    double h2d_ms, sort_ms, link_ms, label_ms, fof6d_ms, potential_ms, unbind_ms, d2h_ms, total_ms;
    long long n_cells, n_groups3d, n_groups_final, n_grouped, n_pairs_potential, n_marginal;
    long long h2d_bytes, d2h_bytes, kernel_launches;
    long long n_tree_interactions, n_tree_nodes; /* kd-tree potential: cell + leaf-particle interactions, nodes built */
    long long n_link_queue;                      /* 3DFOF: dense cell pairs finished by the warp-per-pair kernel */
} vrgpu_stats;

/* Create / destroy a context on CUDA device `device` (-1 = current device). */
int vrgpu_init(int device, vrgpu_ctx **out);
void vrgpu_destroy(vrgpu_ctx *ctx);
const char *vrgpu_last_error(const vrgpu_ctx *ctx);
/* Library / build identification: returns "vrgpu <version> sm_100a". */
const char *vrgpu_version(void);
void vrgpu_get_stats(const vrgpu_ctx *ctx, vrgpu_stats *out);

/* 3D FOF.  Replaces NBody::KDTree::FOF (NBodylib/src/KDTree/KDFOF.cxx:12-129) as called by
 * SearchFullSet (src/search.cxx:203): groups = transitive closure of the strict fp64 predicate
 * d^2 < b*b (KDLeafNode.cxx:556-558), periodic images as KDSplitNode.cxx:1496-1540; groups below
 * `minsize` dropped; ids 1..G by size descending, ties by minimum pid.  pfof_out[i] is indexed by the
 * particle's position in the view. */
int vrgpu_fof3d(vrgpu_ctx *ctx, vrgpu_view view, double b, const double *period /*NULL = open, else [3] equal*/,
                long long minsize, long long *pfof_out, long long *ngroups_out);

/* AdjustStructureForPeriod (src/search.cxx:856-895): writes the unwrapped positions back through a
 * writable view (pos field only).  view.base is cast to non-const. */
int vrgpu_adjust_for_period(vrgpu_ctx *ctx, vrgpu_view view, double period, const long long *pfof, long long ngroups);

/* 6D FOF inside every 3D group.  Replaces src/search.cxx:400-698 (regroup, per-group dispersion,
 * ScalePhase, 6-D KDTree::FOF(1.0), unscale, ReorderGroupIDs).  mode 3 = FOF6DADAPTIVE, 4 = FOF6D.
 * b6 = sqrt(param[1]) after search.cxx:410.  The view must be writable: positions/velocities receive the
 * reference's scale/unscale rounding (two multiplies). */
int vrgpu_fof6d(vrgpu_ctx *ctx, vrgpu_view view, const long long *pfof3d, long long ngroups3d, double b6,
                double vfac, int mode, long long minsize, long long *pfof_out, long long *ngroups_out);

/* Replaces NBody::KDTree::FOFCriterion(FOF6d, params, numgroups, minsize, order = 1, ipcheckflag = 1, FOFcheckbg)
 * (NBodylib/src/KDTree/KDFOF.cxx:133-240) as the 6D core search of SearchSubset calls it (src/search.cxx:1607-1611 and
 * :1664): particles i with check[i] >= check_below are left out (FOFcheckbg, src/fofalgo.cxx:205-208;
 * check == NULL keeps everyone), the others link iff sum_j dx_j^2/xlink2 + dv_j^2/vlink2 < 1 (FOFFunc.h:48-55: params[6],
 * params[7]); groups below minsize dissolve; pfof_out[i] by particle index, ids 1..G by size descending (min pid breaks ties).
 * Open box (the subset is in its halo's centre-of-mass frame). */
int vrgpu_fof_criterion6d(vrgpu_ctx *ctx, vrgpu_view view, const double *check, double check_below, double xlink2, double vlink2,
                          long long minsize, long long *pfof_out, long long *ngroups_out);

/* Local velocity density of the particles of one structure.  Replaces GetVelocityDensity (src/localfield.cxx:18) as
 * SearchSubSub calls it with tree == NULL (src/search.cxx:2646): a physical kd-tree with bucket
 * bsize (Options::Bsize), per leaf (approx = 1, Options::iLocalVelDenApproxCalcFlag, the default; localfield.cxx:726-897) or
 * per particle (approx = 0; :485-586) the nsearch nearest particles in position, of those the nvel nearest in velocity,
 * and KDTree::CalcSmoothLocalValue with the Epanechnikov table of 1000 entries.  type (may be NULL = all 1): Particle::type,
 * particles of type 0 get density 0 (the reference's STRUCDEN builds skip them).  period 0 = open domain; a periodic
 * domain follows the reference's sequence of reflected searches.  density_out[i] by particle index.
 * leaf_offsets (may be NULL; nleaves + 1 entries from 0 to n): for a caller that holds its own kd-tree over the same particles
 * (SearchFullSet passes its FOF tree, src/search.cxx:232) -- the particles are then in that tree's order and leaf l is the index
 * range [leaf_offsets[l], leaf_offsets[l+1]); the per-leaf search uses these leaves instead of rebuilding the tree's. */
int vrgpu_velocity_density(vrgpu_ctx *ctx, vrgpu_view view, const int *type, const long long *leaf_offsets, long long nleaves, double period,
                           int bsize, int nsearch, int nvel, int approx, double *density_out);

/* Potentials of groups given in CSR form (offsets[G+1], index[offsets[G]] = particle index in the view).
 * Replaces the direct-sum loop of Unbind (src/unbind.cxx:398-416) for n <= pp_threshold and
 * Potential() (src/unbind.cxx:1607-1741; kd-tree of KDTree.cxx:438-507, bucket, theta) above it.
 * phi_out[k] corresponds to index[k]. */
int vrgpu_potential(vrgpu_ctx *ctx, vrgpu_view view, long long ngroups, const long long *offsets,
                    const long long *index, double G, double eps, double theta, int bucket,
                    long long pp_threshold, double *phi_out);

/* The particle order of the reference's kd-tree.  Replaces what NBody::KDTree::KDTree / BuildNodes leave in the
 * particle array they are built on (NBodylib/src/KDTree/KDTree.cxx:637-674, 438-507): split dimension = largest spread (first
 * wins), median index start + (size-1)/2 found by MedianPos (:304-337, a quickselect whose swap sequence decides the order --
 * and, among exactly equal coordinates, the membership -- of both halves), leaves of at most `bucket` particles.  Every group
 * of the CSR (as in vrgpu_potential; no empty groups) is a tree of its own: perm_out[offsets[g] + j] = offset inside group g of
 * the member that sits at tree position j.  The potentials (vrgpu_potential, vrgpu_unbind) and the local velocity density use
 * this build; the entry point exists so that parity with the reference's arrangement can be checked directly. */
int vrgpu_tree_order(vrgpu_ctx *ctx, vrgpu_view view, long long ngroups, const long long *offsets, const long long *index, int bucket,
                     long long *perm_out);

/* Binding energies of finished groups.  Replaces GetBindingEnergy
 * (src/substructureproperties.cxx:3884-4255) for the default reference frame (ParticleTypeForRefenceFrame = -1, no
 * per-type mass weights).  The view holds the particles in group order: group g (0-based) = records
 * [offsets[g], offsets[g+1]), offsets[0] = 0 (the reference's numingroup / noffset); gcmvel[3*g..] = pdata[].gcmvel.
 * Groups below large_threshold (ompunbindnum, 1000) use the pair sum (:3922-3940), the others Potential() (:4060-4073).
 * Per particle (may be NULL): phi_out (Particle::SetPotential), E_out (kinetic + potential, what the reference leaves in
 * Particle::SetDensity).  Per group: T, Pot (pair sum for small groups, sum of the members' potentials for large ones --
 * the reference's own convention), Efrac, and the member offsets inside the group of the minimum potential (:4003-4016)
 * and the most bound particle (:4216-4241). */
int vrgpu_binding_energy(vrgpu_ctx *ctx, vrgpu_view view, long long ngroups, const long long *offsets, const double *gcmvel,
                         double G, double eps, double theta, int bucket, long long large_threshold, double *phi_out,
                         double *E_out, double *T_out, double *Pot_out, double *Efrac_out, long long *iminpot_out,
                         long long *ibound_out);

/* Ordering of every group's members by a key.  Replaces the per-group qsort of SortAccordingtoBindingEnergy
 * (src/substructureproperties.cxx:4302-4310; DenCompare = total energy, PotCompare = potential): key[] holds one value per
 * grouped particle in group order (offsets as in vrgpu_binding_energy); order_out[offsets[g] + j] = offset inside group g
 * of the member with the j-th smallest key.  Equal keys keep ascending member order (qsort leaves them in any order). */
int vrgpu_sort_groups_by_key(vrgpu_ctx *ctx, long long ngroups, const long long *offsets, const double *key,
                             long long *order_out);

/* Group sizes and member lists.  Replaces BuildNumInGroup + BuildPGList
 * (src/buildandsortarrays.cxx:11-16, 26-43): numingroup_out[0..ngroups] (entry 0 = 0), index_out = the members of group 1,
 * then group 2, ... each in ascending particle index, i.e. pglist[g] = index_out + numingroup_out[1] + ... +
 * numingroup_out[g-1]; index_out must hold n entries (only *ngrouped_out are written).  pfof == NULL takes the labels the
 * last vrgpu_search_full_set_resident / vrgpu_dist_search_full_set_resident call left on the device (n, ngroups ignored). */
int vrgpu_build_pglist(vrgpu_ctx *ctx, const long long *pfof, long long n, long long ngroups, long long *numingroup_out,
                       long long *index_out, long long *ngrouped_out);

/* CheckUnboundGroups + Unbind (src/unbind.cxx:196-304, 320-729).  pfof in/out (0 = unbound), group ids
 * re-ordered by size when anything changed (ireorder=1).  *changed_out = the reference's return flag. */
int vrgpu_unbind(vrgpu_ctx *ctx, vrgpu_view view, long long *pfof_inout, long long *ngroups_inout,
                 const vrgpu_unbind_opts *opts, int *changed_out, long long *n_marginal_out);

/* The whole of SearchFullSet's DM path in one call with the particles resident on the device between the
 * stages: 3DFOF -> AdjustStructureForPeriod -> [6DFOF] -> [CheckUnboundGroups]  (src/search.cxx:24-854).
 * The view must be writable when period > 0 or fofbgtype <= 4 (positions/velocities are updated as the
 * reference updates them). */
int vrgpu_search_full_set(vrgpu_ctx *ctx, vrgpu_view view, const vrgpu_search_opts *opts,
                          long long *pfof_out, long long *ngroups_out);

/* ---- device-resident variant ---------------------------------------------------------------------------
 * The reference is also driven as a library on particles its caller already holds in memory
 * (InvokeVelociraptor, src/swiftinterface.cxx:271-598, calls SearchFullSet at :419 on the simulation's own
 * particle array).  The equivalent here: hand the particles over ONCE, keep them in HBM, and run the path on
 * them any number of times (each run restarts from the pristine copy, because the path rewrites positions
 * and velocities exactly as the reference does).  Results stay on the device until asked for. */
int vrgpu_particles_upload(vrgpu_ctx *ctx, vrgpu_view view);
int vrgpu_particles_release(vrgpu_ctx *ctx);
int vrgpu_search_full_set_resident(vrgpu_ctx *ctx, const vrgpu_search_opts *opts, long long *ngroups_out);
/* copy the pfof of the last resident run to the host (n entries, caller's particle order) */
int vrgpu_resident_pfof(vrgpu_ctx *ctx, long long *pfof_out);

/* ---- multi-GPU: one process (one context) per GPU, x-slab domain decomposition ------------------------------
 * Replaces the MPI build of SearchFullSet (src/search.cxx:242-347; mpiroutines.cxx MPIInitialDomainDecomposition :49,
 * MPIBuildParticleExportList :867, MPILinkAcross :2387, MPIGroupExchange :2548, MPICompileGroups :2818).
 * Rank r of `nranks` holds the particles with slab_lo <= x < slab_hi; the slabs are ordered by rank and tile
 * [0, period).  Rank 0 creates the 128-byte id, the host code broadcasts it (MPI_Bcast in the reference's world,
 * torch.distributed in bench.py), every rank calls vrgpu_dist_init.  The search returns GLOBAL group ids (identical on
 * every rank and identical to the single-GPU ids); *ngroups_out is the global group count.  Like the single-GPU call it
 * rewrites the caller's (writable) records: grouped particles unwrapped across the period about their group's smallest-pid
 * member (search.cxx:856-895; a rank does not know the caller's global order, and the two agree when pids follow the input
 * order), and after a 6DFOF the scale / unscale rounding of positions and velocities (search.cxx:565-575). */
#define VRGPU_UNIQUE_ID_BYTES 128
int vrgpu_dist_unique_id(char *id_out /*[128]*/);
int vrgpu_dist_init(vrgpu_ctx *ctx, int nranks, int rank, const char *id /*[128]*/);
int vrgpu_dist_search_full_set(vrgpu_ctx *ctx, vrgpu_view local, double slab_lo, double slab_hi, const vrgpu_search_opts *opts,
                               long long *pfof_out, long long *ngroups_out);
/* the same on the particles handed over with vrgpu_particles_upload (result: vrgpu_resident_pfof) */
int vrgpu_dist_search_full_set_resident(vrgpu_ctx *ctx, double slab_lo, double slab_hi, const vrgpu_search_opts *opts,
                                        long long *ngroups_out);

/* ---- allocation-free variant ------------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * Direct-sum potentials of ONE group, phi_j = -G m_j sum_{k != j} m_k / sqrt(r_jk^2 + eps^2) -- the arithmetic and the kernel
 * of vrgpu_potential with pp_threshold < 0 (velociraptor-stf_b200/csrc/potential.cu) -- with every buffer supplied by the
 * caller and the kernel's scratch (group-major SoA columns) in a file-scope device array of the library:
 *   - UNLIKE every other entry point the pointers are DEVICE pointers on `device` (-1 = the current device): pos_dev[3n]
 *     (x, y, z per particle), mass_dev[n] (NULL = unit masses), phi_dev[n] (output).  `stream` is a cudaStream_t (NULL = the
 *     default stream); the kernels are queued on it and the call returns without synchronising.
 *   - n <= VRGPU_NOALLOC_MAX_PARTICLES; a larger n is rejected with VRGPU_EINVAL and a message (use vrgpu_potential).
 *   - the call makes no cudaMalloc / cudaFree / pool allocation and needs no context.  The scratch (2.25 MiB) is held by every
 *     GPU that loads the library until exit.  All calls share it: two calls whose kernels overlap on the GPU (different
 *     streams or threads) race -- serialise them.  Every word a call reads it writes first, so no state carries over.
 *   - the first call per device resolves the array's device address (cudaGetSymbolAddress); make it outside graph capture.
 *   - errbuf (may be NULL) receives the error text, at most errbuf_len bytes with the terminator. */
#define VRGPU_NOALLOC_MAX_PARTICLES 65536
int vrgpu_potential_direct_noalloc(int device, void *stream, const double *pos_dev, const double *mass_dev, long long n, double G,
                                   double eps, double *phi_dev, char *errbuf, long long errbuf_len);

/* ---- measurement hooks ------------------------------------------------------------------------------- */
/* the CUDA stream (cudaStream_t) every kernel of this context is launched on, for event timing by the caller */
void *vrgpu_stream(vrgpu_ctx *ctx);
/* Per-kernel device times accumulated since the last reset: one line per kernel group,
 * "<name> <milliseconds> <launch-brackets>\n".  Returns a pointer owned by the context. */
const char *vrgpu_kernel_times(vrgpu_ctx *ctx);
/* measured fp64 FMA throughput of the device in TFLOP/s (denominator of the potential kernels' roofline) */
int vrgpu_measure_fp64_tflops(vrgpu_ctx *ctx, double *tflops_out);
/* microbenchmark: out4[0] vector DFMA alone, out4[1] tensor-core DMMA (mma.sync m8n8k4 f64) alone, out4[2] / out4[3] the
 * DFMA / DMMA rates when one kernel issues both -- TFLOP/s; tells whether the two share a pipe on this device */
int vrgpu_probe_fp64_pipes(vrgpu_ctx *ctx, double *out4);
void vrgpu_kernel_times_reset(vrgpu_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif
