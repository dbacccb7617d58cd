/* oracle/vr_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the VELOCIraptor structure-finding hot path (3DFOF -> 6DFOF -> unbinding).
 * It exists to CHECK the CUDA path (tests/, __graft_entry__.smoke(), bench.py cpu_baseline); the product
 * library never links, loads or calls it.  Parity status: PINNED against the reference itself compiled
 * into oracle/_ref (tests/test_oracle_vs_ref.py) and against vectors generated from it (tests/golden/).
 *
 * Every function cites the reference code it follows (paths under /root/reference).
 */
#ifndef VR_ORACLE_H
#define VR_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    double G;                 /* Options::G */
    double eps;               /* uinfo.eps (softening, already in code length) */
    double Eratio;            /* uinfo.Eratio */
    double minEfrac;          /* uinfo.minEfrac */
    int unbindtype;           /* 0 USYSANDPART, 1 UPART  (allvars.h:211-213) */
    int bgpot;                /* uinfo.bgpot (1: keep background potential) */
    double fracpotref;        /* uinfo.fracpotref: 1.0 = CM velocity of the whole group (unbind.cxx:440-459), else the frame of the
                                 innermost max(npotref, fracpotref * n) members (unbind.cxx:460-544) */
    double maxunbindfrac;     /* 0.5 */
    double maxunboundfracforiterativeunbind; /* 0.95 */
    double maxallowedunboundfrac;            /* 0.025 */
    double theta;             /* uinfo.TreeThetaOpen 0.5 */
    int bucket;               /* uinfo.BucketSize 8 */
    long long min_size;       /* Options::MinSize at the call */
    long long pp_threshold;   /* UNBINDNUM: groups with n <= this use the direct sum (allvars.h:210); <0: all direct */
    long long large_threshold;/* ompunbindnum 1000 (ompvar.h:26) */
    int sortflag_mode;        /* 0: oldnumingroup initialised per group (intended semantics, what the GPU does)
                                 1: literal serial reference: oldnumingroup carried across groups,
                                    starting from oldn_init (the reference reads it UNINITIALISED: stack garbage) */
    long long oldn_init;      /* mode 1 only: the garbage value assumed for the uninitialised oldnumingroup */
    int cmvelreftype;         /* uinfo.cmvelreftype when fracpotref != 1: 0 CMVELREF (radii about the centre of mass), 1 POTREF
                                 (about the member with the lowest potential) -- allvars.h:216-218 */
    long long npotref;        /* uinfo.Npotref (20) */
} vro_unbind_opts;

/* KDTree::FOF with TPHYS (KDFOF.cxx:12-129, KDLeafNode.cxx:518-573, KDSplitNode.cxx:1496-1540):
 * transitive closure of the strict fp64 predicate d^2 < b*b, periodic images evaluated with the target shifted
 * by +-period.  Groups smaller than minsize are dropped; ids are 1..G by (size desc, min pid asc).
 * pid may be NULL (index is used).  Returns G. */
long long vro_fof3d(long long n, const double *pos, double b, const double *period, long long minsize,
                    const long long *pid, long long *pfof_out);

/* KDTree::FOFCriterion with FOF6d + FOFcheckbg, the linking step of the 6D core search (search.cxx:1607-1611, 1664;
 * SURVEY section 8 row f2): particles with pot[i] >= p9 are left out, the others link iff
 * sum_j dx_j^2/p6 + dv_j^2/p7 < 1 (FOFFunc.h:48-55, that summation order).  Open box.  Returns G. */
long long vro_fof_criterion6d(long long n, const double *pos, const double *vel, const double *pot, double p6, double p7,
                              double p9, long long minsize, const long long *pid, long long *pfof_out);

/* AdjustStructureForPeriod (search.cxx:856-895): pos modified in place. */
void vro_adjust_for_period(long long n, double *pos, double period, const long long *pfof, long long ngroups);

/* 6DFOF inside each 3D group (search.cxx:400-698).  mode 3 = FOF6DADAPTIVE (per-group dispersion),
 * mode 4 = FOF6D (single scale from the largest group, including the reference's quirk Q1).
 * pos/vel are scaled and unscaled in place exactly like the reference (two multiplies, quirk Q3).
 * vscale2_out (may be NULL) receives vscale2array[1..ngroups3d] at index g. Returns number of 6D groups. */
long long vro_fof6d(long long n, double *pos, double *vel, const double *mass, const long long *pfof3d,
                    long long ngroups3d, double b6, double vfac, int mode, long long minsize,
                    const long long *pid, long long *pfof_out, double *vscale2_out);

/* Direct-sum potential of one group in the reference's pair order (unbind.cxx:398-416). */
void vro_potential_pp(long long n, const double *pos, const double *mass, double G, double eps, double *phi_out);

/* kd-tree monopole potential of one group (Potential(), unbind.cxx:1607-1741; tree build KDTree.cxx:438-507).
 * interactions_out (may be NULL) receives the number of cell+particle interactions (F_tree). */
void vro_potential_tree(long long n, const double *pos, const double *mass, double G, double eps, double theta,
                        int bucket, double *phi_out, long long *interactions_out);

/* Tree order and node ranges of the kd-tree vro_potential_tree builds (KDTree.cxx:304-337, 438-507). */
long long vro_tree_order(long long n, const double *pos, int bucket, long long *perm_out, long long *node_out, long long node_cap);

/* Row f4, oracle side only: GetVelocityDensity (localfield.cxx:18) for one structure, approx = iLocalVelDenApproxCalcFlag.
 * type may be NULL (all particles typed).  Returns the number of leaf nodes, < 0 if nsearch / nvel do not fit n. */
long long vro_velocity_density(long long n, const double *pos, const double *vel, const int *type, double period, int bsize,
                               int nsearch, int nvel, int approx, double *den_out);

/* CheckUnboundGroups + Unbind (unbind.cxx:196-304, 320-729) for fracpotref==1.
 * pfof in/out (0 = unbound), returns the new number of groups, ids re-ordered (size desc, min pid asc) when
 * anything changed.  n_marginal_out (may be NULL): number of grouped particles whose |E| <= tol*|phi| at the
 * moment they were last evaluated.  phi_out (may be NULL, length n): potential used for each grouped particle. */
long long vro_unbind(long long n, const double *pos, const double *vel, const double *mass, long long *pfof,
                     long long ngroups, const vro_unbind_opts *opts, const long long *pid, double marginal_tol,
                     long long *n_marginal_out, double *phi_out, int *changed_out);

/* GetBindingEnergy (substructureproperties.cxx:3884-4255) for the default frame (uinfo.cmvelreftype = CMVELREF,
 * ParticleTypeForRefenceFrame = -1, masses stored): SURVEY 8f row f1, the third potential evaluation of every group.
 * Particles are GROUP ORDERED: group g = 1..ngroups owns [off[g], off[g+1]) with off[1] = 0 (the layout
 * SortAccordingtoBindingEnergy hands over, :4279-4290); gcmvel[3*g..] is the kinetic reference velocity GetCM left in
 * pdata[g].gcmvel.  Groups below large_threshold (ompunbindnum, ompvar.h:27) take the pair loop (:3922-3940) -- and
 * their Pot is the sum over PAIRS; larger groups take Potential() and their Pot is the sum over PARTICLES (:4184: twice
 * the pair sum; reproduced as is).  Outputs: phi[n] (Particle::phi), E[n] = T_i + phi_i (Particle::rho, :4037/:4186),
 * per group T, Pot, Efrac (bound fraction), index within the group of the minimum-potential particle (:4003-4016) and of
 * the most bound particle (:4216-4226). */
void vro_binding_energy(long long n, const double *pos, const double *vel, const double *mass, long long ngroups,
                        const long long *off /* ngroups + 2 */, const double *gcmvel /* 3 * (ngroups + 1) */, double G,
                        double eps, double theta, int bucket, long long large_threshold, double *phi_out, double *E_out,
                        double *T_out, double *Pot_out, double *Efrac_out, long long *iminpot_out, long long *ibound_out);

#ifdef __cplusplus
}
#endif
#endif
