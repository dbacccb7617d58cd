/* velociraptor-stf_b200/host/vrgpu_adaptor.cxx -- the reference-side host adaptor.
 *
 * Replacement bodies, with the reference's own signatures (src/proto.h), for the entry points the B200 path takes over:
 *
 *     Int_t *SearchFullSet(Options&, const Int_t, vector<Particle>&, Int_t&)          src/search.cxx:24
 *     int    CheckUnboundGroups(Options, const Int_t, Particle*, Int_t&, Int_t*&, ...)  src/unbind.cxx:196
 *     void   Potential(Options&, Int_t, Particle*)                                      src/unbind.cxx:1607
 *     void   GetBindingEnergy(Options&, const Int_t, Particle*, Int_t, Int_t*&, Int_t*&, PropData*&, Int_t*&)
 *                                                                    src/substructureproperties.cxx:3884  (row f1)
 *     Int_t *BuildNumInGroup(const Int_t, const Int_t, Int_t*)                src/buildandsortarrays.cxx:11  (row f3)
 *     Int_t **BuildPGList(const Int_t, const Int_t, Int_t*, Int_t*)           src/buildandsortarrays.cxx:26  (row f3)
 *     Int_t *FOFCriterion6DCore(Particle*, Int_t, Double_t*, Int_t&, Int_t)      tree->FOFCriterion(FOF6d, ..., FOFcheckbg) of the
 *                        6D core search, src/search.cxx:1611 and :1664 (NBodylib/src/KDTree/KDFOF.cxx:133)      (row f2)
 *     void   GetVelocityDensity(Options&, const Int_t, Particle*, KDTree*)               src/localfield.cxx:18   (row f4)
 *     void   SortGroupParticlesGPU(Options&, Particle*, Int_t, Int_t*, Int_t*, PropData*)
 *                        the per-group qsort inside SortAccordingtoBindingEnergy, substructureproperties.cxx:4302-4316
 *
 * They marshal nothing: libvrgpu reads the caller's NBody::Particle array through a strided view (vrgpu_view).  This
 * file is compiled against the reference's headers (stf.h) by whoever builds the reference with the GPU path; it is
 * plain host C++ on top of the C ABI in include/vrgpu.h.  To avoid duplicate symbols when it is linked NEXT TO the
 * reference's own search.cxx / unbind.cxx (as oracle/Makefile does for the parity test), the functions carry the
 * suffix GPU unless VRGPU_REPLACE_REFERENCE_SYMBOLS is defined (INTEGRATION.md shows both ways).
 *
 * Error behaviour follows the reference: a message on cerr and exit(), e.g. src/search.cxx:503-517.
 */
#include "stf.h"

#include "vrgpu.h"

#include <cstddef>
#include <cstring>

#ifdef VRGPU_REPLACE_REFERENCE_SYMBOLS
#define VRGPU_NAME(x) x
#else
#define VRGPU_NAME(x) x##GPU
#endif

using namespace std;
using namespace NBody;

/* bodies further down that SearchFullSet uses */
Int_t *VRGPU_NAME(BuildNumInGroup)(const Int_t nbodies, const Int_t numgroups, Int_t *pfof);
void VRGPU_NAME(GetVelocityDensity)(Options &opt, const Int_t nbodies, Particle *Part, KDTree *tree);

static_assert(sizeof(Int_t) == sizeof(long long), "the GPU path needs the default LONGINT build (Int_t == long long)");
static_assert(sizeof(Double_t) == sizeof(double), "the GPU path needs the default double precision build");

namespace {

vrgpu_ctx *g_ctx = NULL;

[[noreturn]] void die(const char *what, int rc)
{
    cerr << "vrgpu: " << what << " failed (" << rc << "): " << (g_ctx ? vrgpu_last_error(g_ctx) : "no context") << endl;
    exit(9);
}

vrgpu_ctx *context()
{
    if (g_ctx == NULL) {
        int rc = vrgpu_init(-1, &g_ctx);
        if (rc != VRGPU_OK) {
            cerr << "vrgpu: no usable B200 / CUDA device (code " << rc << "); there is no CPU fallback in this build" << endl;
            exit(9);
        }
    }
    return g_ctx;
}

/* Field offsets of NBody::Particle, found by probing one object (mass and pid have no address-returning accessor). */
struct Layout { size_t pos, vel, mass, pid; };
const Layout &layout()
{
    static Layout L;
    static bool done = false;
    if (done) return L;
    Particle p;
    memset((void *)&p, 0, sizeof(Particle));
    const unsigned char *base = reinterpret_cast<const unsigned char *>(&p);
    L.pos = reinterpret_cast<const unsigned char *>(p.GetPosition()) - base;
    L.vel = reinterpret_cast<const unsigned char *>(p.GetVelocity()) - base;
    const double magic = 1.2345678901234567e77;
    p.SetMass(magic);
    p.SetPID((PARTPIDTYPE)0x5a5a5a5a1234567LL);
    L.mass = L.pid = (size_t)-1;
    for (size_t o = 0; o + 8 <= sizeof(Particle); o += 8) {
        double d;
        long long q;
        memcpy(&d, base + o, 8);
        memcpy(&q, base + o, 8);
        if (d == magic) L.mass = o;
        if (q == 0x5a5a5a5a1234567LL) L.pid = o;
    }
    if (L.mass == (size_t)-1 || L.pid == (size_t)-1 || sizeof(Particle) % 8 != 0) {
        cerr << "vrgpu: unsupported NBody::Particle layout (SINGLEPRECISION / NOMASS / LOWPRECISIONPOS builds are not supported)" << endl;
        exit(9);
    }
    done = true;
    return L;
}

vrgpu_view view_of(const Particle *Part, Int_t n)
{
    const Layout &L = layout();
    vrgpu_view v;
    v.base = Part;
    v.stride = sizeof(Particle);
    v.off_pos = L.pos; v.off_vel = L.vel; v.off_mass = L.mass; v.off_pid = L.pid;
    v.n = n;
    return v;
}

vrgpu_unbind_opts unbind_opts_of(const Options &opt)
{
    vrgpu_unbind_opts u;
    memset(&u, 0, sizeof u);
    u.G = opt.G;
    u.eps = opt.uinfo.eps;
    u.Eratio = opt.uinfo.Eratio;
    u.minEfrac = opt.uinfo.minEfrac;
    u.unbindtype = opt.uinfo.unbindtype;
    u.bgpot = opt.uinfo.bgpot;
    u.fracpotref = opt.uinfo.fracpotref;
    u.maxunbindfrac = opt.uinfo.maxunbindfrac;
    u.maxunboundfracforiterativeunbind = opt.uinfo.maxunboundfracforiterativeunbind;
    u.maxallowedunboundfrac = opt.uinfo.maxallowedunboundfrac;
    u.theta = opt.uinfo.TreeThetaOpen;
    u.bucket = opt.uinfo.BucketSize;
    u.min_size = opt.MinSize;
    u.pp_threshold = UNBINDNUM;   /* allvars.h:210 */
    u.large_threshold = 1000;     /* ompunbindnum, ompvar.h:26 */
    u.marginal_tol = 1e-6;
    u.cmvelreftype = opt.uinfo.cmvelreftype;
    u.npotref = opt.uinfo.Npotref;
    return u;
}

} // namespace

/* SearchFullSet, DM-only path (src/search.cxx:24-854).  Same contract: returns new Int_t[nbodies] indexed by the
 * particle's array index (caller delete[]s it, main.cxx:427), ids 1..numgroups by size, fills level 0 of the global
 * psldata (search.cxx:64,740-760), leaves the particles in input order with grouped ones unwrapped across the period. */
Int_t *VRGPU_NAME(SearchFullSet)(Options &opt, const Int_t nbodies, vector<Particle> &Part, Int_t &numgroups)
{
    if (opt.partsearchtype != PSTALL && opt.partsearchtype != PSTDARK) {
        cerr << "vrgpu: only Particle_search_type all / dark matter is implemented on the GPU path" << endl;
        exit(9);
    }
    if (opt.iKeepFOF && opt.fofbgtype <= FOF6D) {
        cerr << "vrgpu: Keep_FOF with a 6DFOF field search is not implemented on the GPU path" << endl;
        exit(9);
    }
    if (!(opt.fofbgtype == FOF3D || opt.fofbgtype == FOF6D || opt.fofbgtype == FOF6DADAPTIVE)) {
        cerr << "vrgpu: FoF_Field_search_type must be 5 (3DFOF), 4 (6DFOF) or 3 (adaptive 6DFOF) on the GPU path" << endl;
        exit(9);
    }
    vrgpu_search_opts so;
    memset(&so, 0, sizeof so);
    so.period = opt.p;
    so.param1 = (opt.ellxscale * opt.ellxscale) * (opt.ellphys * opt.ellphys) * (opt.ellhalophysfac * opt.ellhalophysfac); /* search.cxx:77 */
    so.halo_min_size = opt.HaloMinSize;
    so.fofbgtype = opt.fofbgtype;
    so.ellhalo6dxfac = opt.ellhalo6dxfac;
    so.ellhalo6dvfac = opt.ellhalo6dvfac;
    so.bound_halos = opt.iBoundHalos;
    so.unbind = unbind_opts_of(opt);
    Int_t *pfof = new Int_t[nbodies > 0 ? nbodies : 1];
    long long ng = 0;
    int rc;
#ifdef STRUCDEN
    /* search.cxx:205-235: a STRUCDEN build that goes on to the substructure search computes the local velocity density of the
     * members of 3D groups of at least MINSUBSIZE particles here -- on the coordinates as read (before AdjustStructureForPeriod),
     * with the tree of the 3DFOF, which is the tree GetVelocityDensity builds for tree == NULL (same bucket, same particles).
     * So: the 3D labels first (no coordinate is touched), the density with the types set as the reference sets them, then the
     * whole path. */
    if (opt.iSubSearch == 1 && opt.foftype != FOF6DCORE && nbodies > 0) {
        double per[3] = {(double)opt.p, (double)opt.p, (double)opt.p};
        long long ng3 = 0;
        rc = vrgpu_fof3d(context(), view_of(Part.data(), nbodies), sqrt(so.param1), opt.p > 0 ? per : NULL, so.halo_min_size,
                         reinterpret_cast<long long *>(pfof), &ng3);
        if (rc != VRGPU_OK) die("vrgpu_fof3d", rc);
        if (ng3 > 0) {
            Int_t *num3 = VRGPU_NAME(BuildNumInGroup)(nbodies, ng3, pfof);
            vector<int> storetype((size_t)nbodies);
            Int_t numinstrucs = 0;
            for (Int_t i = 0; i < nbodies; i++) {
                storetype[i] = Part[i].GetType();
                int t;
                if (!(opt.iBaryonSearch >= 1 && opt.partsearchtype == PSTALL)) t = num3[pfof[i]] >= MINSUBSIZE;
                else t = storetype[i] == DARKTYPE ? (int)(num3[pfof[i]] >= MINSUBSIZE) : -1;
                Part[i].SetType(t);
                numinstrucs += t > 0;
            }
            if (numinstrucs > 0) VRGPU_NAME(GetVelocityDensity)(opt, nbodies, Part.data(), NULL);
            for (Int_t i = 0; i < nbodies; i++) Part[i].SetType(storetype[i]);
            delete[] num3;
        }
    }
#endif
    rc = vrgpu_search_full_set(context(), view_of(Part.data(), nbodies), &so, reinterpret_cast<long long *>(pfof), &ng);
    if (rc != VRGPU_OK) die("vrgpu_search_full_set", rc);
    numgroups = ng;
    /* the reference rewrites the scratch id field to the array index (KDTree construction, search.cxx:692-693) */
    for (Int_t i = 0; i < nbodies; i++) Part[i].SetID(i);
    /* hierarchy level 0 (search.cxx:64, 740-760) */
    psldata = new StrucLevelData;
    psldata->Allocate(numgroups);
    psldata->Initialize();
    for (Int_t i = 0; i < nbodies; i++) {
        if (pfof[i] > 0 && psldata->gidhead[pfof[i]] == NULL) {
            psldata->gidhead[pfof[i]] = &pfof[i];
            psldata->Phead[pfof[i]] = &Part[i];
            psldata->gidparenthead[pfof[i]] = &pfof[i];
            psldata->giduberparenthead[pfof[i]] = &pfof[i];
            psldata->stypeinlevel[pfof[i]] = HALOSTYPE;
        }
    }
    psldata->stype = HALOSTYPE;
    return pfof;
}

/* CheckUnboundGroups (src/unbind.cxx:196-304): pfof updated in place (0 = unbound), ngroup = the number of groups that survive,
 * optional numingroup / pglist refreshed with the surviving members, returns 1 if anything changed.  The originals in Part are
 * untouched, as in the reference (it works on copies).  The three forms the reference calls:
 *   ireorder = 1, groupflag = NULL (search.cxx:703): ids compacted and re-ordered by size (the library's canonical order);
 *   ireorder = 0 (search.cxx:2867, :3452): every surviving group KEEPS its id, dissolved groups have numingroup = 0
 *     (unbind.cxx:718-720; groupflag, if given, is not touched on this branch);
 *   ireorder = 1 with groupflag (search.cxx:2674, the core flags of SearchSubSub): ids re-ordered by the reference's group value
 *     -- flagged groups (flag == 1) first by size descending, the others by 1/size, i.e. size ascending -- and groupflag
 *     permuted with them, dissolved groups get flag 0 (unbind.cxx:271-287; ReorderGroupIDsAndArraybyValue,
 *     buildandsortarrays.cxx:238-253).  Equal values: the smaller minimum particle index first (the reference's heap order
 *     among equal values is not a function of the input, quirk Q5). */
int VRGPU_NAME(CheckUnboundGroups)(Options opt, const Int_t nbodies, Particle *Part, Int_t &ngroup, Int_t *&pfof, Int_t *numingroup,
                                   Int_t **pglist, int ireorder, Int_t *groupflag)
{
    vrgpu_unbind_opts u = unbind_opts_of(opt);
    const Int_t ngroup_in = ngroup;
    const bool keep_ids = ireorder != 1 || groupflag != NULL;
    Int_t *oldpf = NULL;
    if (keep_ids) {
        oldpf = new Int_t[nbodies > 0 ? nbodies : 1];
        for (Int_t i = 0; i < nbodies; i++) oldpf[i] = pfof[i];
    }
    long long ng = ngroup;
    int changed = 0;
    long long nmarginal = 0;
    int rc = vrgpu_unbind(context(), view_of(Part, nbodies), reinterpret_cast<long long *>(pfof), &ng, &u, &changed, &nmarginal);
    if (rc != VRGPU_OK) die("vrgpu_unbind", rc);
    ngroup = ng;
    if (keep_ids) {
        /* unbinding neither merges nor splits groups: a surviving particle's old id names its group */
        for (Int_t i = 0; i < nbodies; i++) pfof[i] = pfof[i] > 0 ? oldpf[i] : 0;
        delete[] oldpf;
        if (ireorder == 1 && changed && ngroup > 0) {
            /* sizes, first members and values by OLD id; new id = rank by (value descending, first member ascending) */
            Int_t *num = new Int_t[ngroup_in + 1], *first = new Int_t[ngroup_in + 1], *order = new Int_t[ngroup_in + 1], *newid = new Int_t[ngroup_in + 1];
            Int_t *gtemp = new Int_t[ngroup_in + 1];
            double *value = new double[ngroup_in + 1];
            for (Int_t g = 0; g <= ngroup_in; g++) { num[g] = 0; first[g] = nbodies; newid[g] = 0; }
            for (Int_t i = 0; i < nbodies; i++) if (pfof[i] > 0) { if (num[pfof[i]]++ == 0) first[pfof[i]] = i; }
            Int_t nlive = 0;
            for (Int_t g = 1; g <= ngroup_in; g++) {
                gtemp[g] = groupflag[g];
                if (num[g] == 0) { value[g] = 0; groupflag[g] = 0; }
                else { value[g] = groupflag[g] == 1 ? (double)num[g] : 1.0 / (double)num[g]; order[nlive++] = g; }
            }
            std::sort(order, order + nlive, [&](Int_t a, Int_t b) { return value[a] != value[b] ? value[a] > value[b] : first[a] < first[b]; });
            for (Int_t k = 0; k < nlive; k++) { newid[order[k]] = k + 1; groupflag[k + 1] = gtemp[order[k]]; }
            for (Int_t i = 0; i < nbodies; i++) if (pfof[i] > 0) pfof[i] = newid[pfof[i]];
            delete[] num; delete[] first; delete[] order; delete[] newid; delete[] gtemp; delete[] value;
        }
    }
    if (numingroup != NULL) {
        for (Int_t g = 0; g <= ngroup_in; g++) numingroup[g] = 0;
        for (Int_t i = 0; i < nbodies; i++) if (pfof[i] > 0) numingroup[pfof[i]]++;
    }
    if (pglist != NULL && numingroup != NULL) {
        /* lists are re-filled with the surviving indices (unbind.cxx:142-144); the caller's arrays are large enough
         * because groups only shrink -- after a re-ordering the caller rebuilds them (search.cxx:2675-2682, unbind.cxx:296) */
        const bool same_ids = ireorder != 1 || !changed;
        if (same_ids) {
            Int_t *fill = new Int_t[ngroup_in + 1];
            for (Int_t g = 0; g <= ngroup_in; g++) fill[g] = 0;
            for (Int_t i = 0; i < nbodies; i++) if (pfof[i] > 0) pglist[pfof[i]][fill[pfof[i]]++] = i;
            delete[] fill;
        }
    }
    return changed;
}

/* Potential (src/unbind.cxx:1607-1741): kd-tree monopole potential of ONE set of particles, written to Part[j].phi. */
void VRGPU_NAME(Potential)(Options &opt, Int_t nbodies, Particle *Part)
{
    if (nbodies <= 0) return;
    long long offsets[2] = {0, nbodies};
    long long *index = new long long[nbodies];
    double *phi = new double[nbodies];
    for (Int_t i = 0; i < nbodies; i++) index[i] = i;
    /* Potential() always uses the tree: pp_threshold 0 sends every group with n > 0 to it */
    int rc = vrgpu_potential(context(), view_of(Part, nbodies), 1, offsets, index, opt.G, opt.uinfo.eps, opt.uinfo.TreeThetaOpen,
                             opt.uinfo.BucketSize, 0, phi);
    if (rc != VRGPU_OK) die("vrgpu_potential", rc);
    for (Int_t i = 0; i < nbodies; i++) Part[i].SetPotential(phi[i]);
    delete[] phi;
    delete[] index;
}

/* GetBindingEnergy (src/substructureproperties.cxx:3884-4255), default frame.  Part is in group order (numingroup /
 * noffset), pdata[].gcmvel holds the frame GetCM left.  Writes what the reference writes: Part[].phi (SetPotential),
 * the total energy in Part[].rho (SetDensity), pdata[].T / Pot / Efrac, iminpot + gposminpot / gvelminpot, ibound +
 * gposmbp / gvelmbp.  Builds with GASON / STARON / NOMASS, the minimum-potential frame (cmvelreftype = POTREF), a
 * restricted ParticleTypeForRefenceFrame and pre-loaded potentials (icalculatepotential = 0) stay on the CPU path. */
void VRGPU_NAME(GetBindingEnergy)(Options &opt, const Int_t nbodies, Particle *Part, Int_t ngroup, Int_t *&pfof, Int_t *&numingroup,
                                  PropData *&pdata, Int_t *&noffset)
{
#if defined(GASON) || defined(STARON) || defined(NOMASS)
#error "the GPU GetBindingEnergy covers the dark-matter build only (no GASON / STARON / NOMASS)"
#endif
    (void)pfof;
    if (opt.uinfo.cmvelreftype == POTREF || opt.ParticleTypeForRefenceFrame != -1 || !opt.uinfo.icalculatepotential) {
        cerr << "vrgpu: GetBindingEnergy on the GPU path needs the CM frame, all particle types and Calculate potential on" << endl;
        exit(9);
    }
    if (ngroup <= 0) return;
    long long *offsets = new long long[ngroup + 1];
    double *gcmvel = new double[3 * ngroup];
    Int_t ntot = 0;
    for (Int_t g = 1; g <= ngroup; g++) {
        if (noffset[g] != ntot) {
            cerr << "vrgpu: GetBindingEnergy expects particles sorted by group (noffset[" << g << "] = " << noffset[g] << ", expected " << ntot << ")" << endl;
            exit(9);
        }
        offsets[g - 1] = ntot;
        ntot += numingroup[g];
        for (int k = 0; k < 3; k++) gcmvel[3 * (g - 1) + k] = pdata[g].gcmvel[k];
    }
    offsets[ngroup] = ntot;
    if (ntot > nbodies) {
        cerr << "vrgpu: GetBindingEnergy: numingroup sums past nbodies" << endl;
        exit(9);
    }
    double *phi = new double[ntot > 0 ? ntot : 1], *E = new double[ntot > 0 ? ntot : 1];
    double *T = new double[ngroup], *Pot = new double[ngroup], *Efrac = new double[ngroup];
    long long *iminpot = new long long[ngroup], *ibound = new long long[ngroup];
    int rc = vrgpu_binding_energy(context(), view_of(Part, ntot), ngroup, offsets, gcmvel, opt.G, opt.uinfo.eps, opt.uinfo.TreeThetaOpen,
                                  opt.uinfo.BucketSize, ompunbindnum, phi, E, T, Pot, Efrac, iminpot, ibound);
    if (rc != VRGPU_OK) die("vrgpu_binding_energy", rc);
    for (Int_t i = 0; i < ntot; i++) {
        Part[i].SetPotential(phi[i]);
        Part[i].SetDensity(E[i]);
    }
    for (Int_t g = 1; g <= ngroup; g++) {
        if (numingroup[g] <= 0) continue;
        pdata[g].T = T[g - 1];
        pdata[g].Pot = Pot[g - 1];
        pdata[g].Efrac = Efrac[g - 1];
        const Particle &pmin = Part[noffset[g] + iminpot[g - 1]], &pmb = Part[noffset[g] + ibound[g - 1]];
        pdata[g].iminpot = pmin.GetPID();
        pdata[g].ibound = pmb.GetPID();
        for (int k = 0; k < 3; k++) {
            pdata[g].gposminpot[k] = pmin.GetPosition(k);
            pdata[g].gvelminpot[k] = pmin.GetVelocity(k);
            pdata[g].gposmbp[k] = pmb.GetPosition(k);
            pdata[g].gvelmbp[k] = pmb.GetVelocity(k);
        }
    }
    delete[] ibound; delete[] iminpot; delete[] Efrac; delete[] Pot; delete[] T; delete[] E; delete[] phi;
    delete[] gcmvel; delete[] offsets;
}

/* The per-group sort of SortAccordingtoBindingEnergy (src/substructureproperties.cxx:4302-4316): every group's particles
 * re-arranged in ascending total energy (Sort_by_binding_energy = 1: the value GetBindingEnergy left in
 * Particle::GetDensity, DenCompare) or ascending potential (PotCompare); pdata[].iunbound = position of the first
 * member with positive energy, the group size if there is none.  The keys go to the device, the ordering comes back and
 * the 96-byte records are permuted here, group by group.  Equal keys keep their order (qsort leaves them in any order). */
void SortGroupParticlesGPU(Options &opt, Particle *Part, Int_t ngroup, Int_t *numingroup, Int_t *noffset, PropData *pdata)
{
    if (ngroup <= 0) return;
    long long *offsets = new long long[ngroup + 1];
    Int_t ntot = 0, nmax = 0;
    for (Int_t g = 1; g <= ngroup; g++) {
        if (noffset[g] != ntot) {
            cerr << "vrgpu: SortGroupParticles expects particles sorted by group (noffset[" << g << "] = " << noffset[g] << ", expected " << ntot << ")" << endl;
            exit(9);
        }
        offsets[g - 1] = ntot;
        ntot += numingroup[g];
        if (numingroup[g] > nmax) nmax = numingroup[g];
    }
    offsets[ngroup] = ntot;
    double *key = new double[ntot > 0 ? ntot : 1];
    long long *order = new long long[ntot > 0 ? ntot : 1];
    for (Int_t i = 0; i < ntot; i++) key[i] = opt.iSortByBindingEnergy ? Part[i].GetDensity() : Part[i].GetPotential();
    int rc = vrgpu_sort_groups_by_key(context(), ngroup, offsets, key, order);
    if (rc != VRGPU_OK) die("vrgpu_sort_groups_by_key", rc);
    vector<Particle> tmp(nmax > 0 ? nmax : 1);
    for (Int_t g = 1; g <= ngroup; g++) {
        const Int_t n = numingroup[g], o = noffset[g];
        for (Int_t j = 0; j < n; j++) tmp[j] = Part[o + j];
        for (Int_t j = 0; j < n; j++) Part[o + j] = tmp[order[o + j]];
        pdata[g].iunbound = n;
        for (Int_t j = 0; j < n; j++) if (Part[o + j].GetDensity() > 0) { pdata[g].iunbound = j; break; }
    }
    delete[] order;
    delete[] key;
    delete[] offsets;
}

/* BuildNumInGroup (src/buildandsortarrays.cxx:11-16): new Int_t[numgroups + 1], entry 0 = 0.  When pfof is the array the
 * last SearchFullSet returned, a caller that also needs the lists should call BuildPGList below, which refreshes the
 * sizes in the same device pass. */
Int_t *VRGPU_NAME(BuildNumInGroup)(const Int_t nbodies, const Int_t numgroups, Int_t *pfof)
{
    Int_t *numingroup = new Int_t[numgroups + 1];
    Int_t *index = new Int_t[nbodies > 0 ? nbodies : 1];
    long long ngrouped = 0;
    int rc = vrgpu_build_pglist(context(), reinterpret_cast<long long *>(pfof), nbodies, numgroups, reinterpret_cast<long long *>(numingroup),
                                reinterpret_cast<long long *>(index), &ngrouped);
    if (rc != VRGPU_OK) die("vrgpu_build_pglist", rc);
    delete[] index;
    return numingroup;
}

/* BuildPGList (src/buildandsortarrays.cxx:26-43): pglist[g] = new Int_t[numingroup[g]] holding the members of group g in
 * ascending particle index, NULL for empty groups; numingroup is re-counted, as the reference does. */
Int_t **VRGPU_NAME(BuildPGList)(const Int_t nbodies, const Int_t numgroups, Int_t *numingroup, Int_t *pfof)
{
    Int_t **pglist = new Int_t *[numgroups + 1];
    Int_t *index = new Int_t[nbodies > 0 ? nbodies : 1];
    long long ngrouped = 0;
    int rc = vrgpu_build_pglist(context(), reinterpret_cast<long long *>(pfof), nbodies, numgroups, reinterpret_cast<long long *>(numingroup),
                                reinterpret_cast<long long *>(index), &ngrouped);
    if (rc != VRGPU_OK) die("vrgpu_build_pglist", rc);
    Int_t at = 0;
    for (Int_t g = 1; g <= numgroups; g++) {
        pglist[g] = NULL;
        if (numingroup[g] <= 0) continue;
        pglist[g] = new Int_t[numingroup[g]];
        memcpy(pglist[g], index + at, sizeof(Int_t) * (size_t)numingroup[g]);
        at += numingroup[g];
    }
    delete[] index;
    return pglist;
}

#ifndef VRGPU_REPLACE_REFERENCE_SYMBOLS
/* A build that routes the call sites by renaming (-DBuildPGList=BuildPGListGPU, oracle/Makefile's stf_gpu) renames every
 * overload.  The two index-indirect ones (src/buildandsortarrays.cxx:63-81, 83-100: lists of Part[i].GetID() / ids[i], used by
 * SearchSubset on its sub-volumes) are not on the path: they forward to the reference's own bodies. */
Int_t **BuildPGListGPU(const Int_t nbodies, const Int_t numgroups, Int_t *numingroup, Int_t *pfof, Particle *Part)
{
    return BuildPGList(nbodies, numgroups, numingroup, pfof, Part);
}
Int_t **BuildPGListGPU(const Int_t nbodies, const Int_t numgroups, Int_t *numingroup, Int_t *pfof, Int_t *ids)
{
    return BuildPGList(nbodies, numgroups, numingroup, pfof, ids);
}
#endif

/* Row f2, linking step.  The 6D core search of SearchSubset (src/search.cxx:1531-1700) calls
 *     pfofbg = tree->FOFCriterion(fofcmp = FOF6d, param, numgroupsbg, minsize, iorder = 1, icheck = 1, FOFcheckbg);
 * at :1611 and, inside its loop over shrinking linking lengths, at :1664.  This body takes the same particles (in whatever
 * order the tree left them), the same param[] (param[6], param[7]: squared linking lengths, param[9]: the FOFcheckbg
 * threshold on the particles' potential field) and returns the same thing: a new Int_t[nsubset] indexed by particle ID
 * (KDFOF.cxx:153-156), 0 = no group, ids 1..numgroups by size descending. */
Int_t *VRGPU_NAME(FOFCriterion6DCore)(Particle *Partsubset, Int_t nsubset, Double_t *param, Int_t &numgroups, Int_t minsize)
{
    Int_t *pfof = new Int_t[nsubset > 0 ? nsubset : 1];
    numgroups = 0;
    if (nsubset <= 0) return pfof;
    vector<double> check((size_t)nsubset);
    for (Int_t i = 0; i < nsubset; i++) check[i] = Partsubset[i].GetPotential();
    vector<long long> byindex((size_t)nsubset);
    long long ng = 0;
    int rc = vrgpu_fof_criterion6d(context(), view_of(Partsubset, nsubset), check.data(), param[9], param[6], param[7], minsize, byindex.data(),
                                   &ng);
    if (rc != VRGPU_OK) die("vrgpu_fof_criterion6d", rc);
    for (Int_t i = 0; i < nsubset; i++) pfof[Partsubset[i].GetID()] = byindex[i];
    numgroups = ng;
    return pfof;
}

/* Row f4.  GetVelocityDensity (src/localfield.cxx:18): Particle::density of every particle (of every particle with type != 0
 * in STRUCDEN builds) from Options::Nsearch position neighbours and Options::Nvel velocity neighbours, per leaf of the kd-tree
 * with bucket Options::Bsize (iLocalVelDenApproxCalcFlag, the default) or per particle.  The library builds the tree the
 * routine would build for tree == NULL (the call of SearchSubSub, search.cxx:2646).  From a tree handed in by the caller
 * (search.cxx:232: the FOF tree over the same particles, which are then in that tree's order) the body reads the leaf
 * boundaries the way the reference itself does (FindLeafNode, localfield.cxx:772-781) and passes them on, so the per-leaf
 * search runs on the caller's leaves.  The smoothing kernel and the periodicity are the ones the routine's own tree has
 * (Epanechnikov table of 1000 entries, period = Options::p), which is also what the tree of search.cxx:232 has; a caller's
 * tree with other settings -- the MPI branch's tree of search.cxx:383 is built with a table of 100 entries, the OpenMP one of
 * :88 without a period -- is outside what this body reproduces (KDTree offers no way to ask: GetPeriod() dereferences a null
 * pointer on an open tree, the kernel resolution is private).  Not covered either: the MPI import / export of neighbours (single task only), HALOONLYDEN builds (another reference routine) and
 * the baryon search variant that restricts neighbours by type (iBaryonSearch == 1 with PSTALL). */
#ifdef HALOONLYDEN
#error "vrgpu_adaptor.cxx: HALOONLYDEN builds use GetVelocityDensityHaloOnlyDen, which the GPU path does not replace"
#endif
void VRGPU_NAME(GetVelocityDensity)(Options &opt, const Int_t nbodies, Particle *Part, KDTree *tree)
{
    if (nbodies <= 0) return;
    if (opt.iBaryonSearch == 1 && opt.partsearchtype == PSTALL) {
        cerr << "vrgpu: GetVelocityDensity with a type-restricted neighbour search (Baryon_searchflag=1, all particle types) is not supported" << endl;
        exit(9);
    }
    vector<double> den((size_t)nbodies);
    const int *tp = NULL;
#ifdef STRUCDEN
    vector<int> type((size_t)nbodies);
    for (Int_t i = 0; i < nbodies; i++) type[i] = Part[i].GetType();
    tp = type.data();
#endif
    vector<long long> leafoff;
    if (tree != NULL && opt.iLocalVelDenApproxCalcFlag) {
        leafoff.push_back(0);
        for (Int_t ipart = 0; ipart < nbodies;) {
            ipart += tree->FindLeafNode(ipart)->GetCount();
            leafoff.push_back(ipart);
        }
    }
    int rc = vrgpu_velocity_density(context(), view_of(Part, nbodies), tp, leafoff.empty() ? NULL : leafoff.data(),
                                    leafoff.empty() ? 0 : (long long)leafoff.size() - 1, opt.p > 0 ? (double)opt.p : 0.0, (int)opt.Bsize,
                                    (int)opt.Nsearch, (int)opt.Nvel, opt.iLocalVelDenApproxCalcFlag ? 1 : 0, den.data());
    if (rc != VRGPU_OK) die("vrgpu_velocity_density", rc);
    for (Int_t i = 0; i < nbodies; i++)
        if (!tp || tp[i] != 0) Part[i].SetDensity(den[i]);
}
