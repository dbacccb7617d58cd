/* oracle/ref_driver.cxx  --  TEST INFRASTRUCTURE ONLY.
 *
 * A thin extern "C" bridge that lets the parity tests (ctypes) and bench.py's reference arm call the
 * UNMODIFIED VELOCIraptor reference functions that the GPU path replaces:
 *
 *   vrref_kdfof            -> NBody::KDTree::FOF                 (NBodylib/src/KDTree/KDFOF.cxx:12)
 *   vrref_search_full_set  -> SearchFullSet                       (src/search.cxx:24)
 *   vrref_potential        -> Potential(Options&,Int_t,Particle*) (src/unbind.cxx:1607)
 *   vrref_unbind           -> CheckUnboundGroups                  (src/unbind.cxx:196)
 *   vrref_binding_energy   -> GetBindingEnergy                    (src/substructureproperties.cxx:3884)
 *   vrref_sort_by_binding_energy -> SortAccordingtoBindingEnergy   (src/substructureproperties.cxx:4256)
 *   vrref_build_pglist     -> BuildNumInGroup + BuildPGList        (src/buildandsortarrays.cxx:11, 26)
 *   vrref_fofcriterion6d   -> NBody::KDTree::FOFCriterion(FOF6d, ..., FOFcheckbg) (NBodylib/src/KDTree/KDFOF.cxx:133)
 *   vrref_velocity_density -> GetVelocityDensity                   (src/localfield.cxx:18; row f4, oracle side only)
 *
 * It is compiled against the reference headers where they lie and linked with the reference's own
 * objects into oracle/_ref/libvrref*.so (see oracle/Makefile).  Nothing here re-implements the
 * algorithms; it only marshals flat arrays into NBody::Particle and back.
 */
#include "stf.h"
#include <chrono>

#ifdef VRREF_USE_GPU
/* libvrref_gpu.so: the SAME bridge, but the three entry points resolve to the B200 host adaptor
 * (velociraptor-stf_b200/host/vrgpu_adaptor.cxx) instead of the reference's own bodies -- the drop-in exercised through the
 * reference's real Options / NBody::Particle / StrucLevelData types. */
Int_t *SearchFullSetGPU(Options &opt, const Int_t nbodies, vector<NBody::Particle> &Part, Int_t &numgroups);
int CheckUnboundGroupsGPU(Options opt, const Int_t nbodies, NBody::Particle *Part, Int_t &ngroup, Int_t *&pfof, Int_t *numingroup = NULL,
                          Int_t **pglist = NULL, int ireorder = 1, Int_t *groupflag = NULL);
void PotentialGPU(Options &opt, Int_t nbodies, NBody::Particle *Part);
void GetBindingEnergyGPU(Options &opt, const Int_t nbodies, NBody::Particle *Part, Int_t ngroup, Int_t *&pfof, Int_t *&numingroup,
                         PropData *&pdata, Int_t *&noffset);
void SortGroupParticlesGPU(Options &opt, NBody::Particle *Part, Int_t ngroup, Int_t *numingroup, Int_t *noffset, PropData *pdata);
Int_t *BuildNumInGroupGPU(const Int_t nbodies, const Int_t numgroups, Int_t *pfof);
Int_t **BuildPGListGPU(const Int_t nbodies, const Int_t numgroups, Int_t *numingroup, Int_t *pfof);
Int_t *FOFCriterion6DCoreGPU(NBody::Particle *Partsubset, Int_t nsubset, Double_t *param, Int_t &numgroups, Int_t minsize);
void GetVelocityDensityGPU(Options &opt, const Int_t nbodies, NBody::Particle *Part, NBody::KDTree *tree);
#define GetVelocityDensity GetVelocityDensityGPU
#define BuildNumInGroup BuildNumInGroupGPU
#define BuildPGList BuildPGListGPU
#define GetBindingEnergy GetBindingEnergyGPU
#define SearchFullSet SearchFullSetGPU
#define CheckUnboundGroups CheckUnboundGroupsGPU
#define Potential PotentialGPU
#endif

using namespace std;
using namespace NBody;

extern "C" {

/* Options subset that steers the hot path.  Field meaning = the reference Options member cited. */
typedef struct {
    double period;            /* Options::p (0 = open box)                         allvars.h */
    double ellxscale;         /* Options::ellxscale  (LN for cosmological input)  gadgetio.cxx:1417 */
    double ellphys;           /* Options::ellphys                                  */
    double ellhalophysfac;    /* Options::ellhalophysfac (ui.cxx:901-906)          */
    double ellhalo6dxfac;     /* Options::ellhalo6dxfac                            */
    double ellhalo6dvfac;     /* Options::ellhalo6dvfac                            */
    long long halo_min_size;  /* Options::HaloMinSize                              */
    long long min_size;       /* Options::MinSize                                  */
    int fofbgtype;            /* Options::fofbgtype: 5 FOF3D, 4 FOF6D, 3 FOF6DADAPTIVE (allvars.h) */
    int bound_halos;          /* Options::iBoundHalos                              */
    int omp_fof;              /* Options::iopenmpfof                               */
    long long omp_fof_size;   /* Options::openmpfofsize (<=0: keep default)        */
    /* UnbindInfo */
    double G;                 /* Options::G                                        */
    double eps;               /* uinfo.eps (already multiplied by LN)              */
    double Eratio;            /* uinfo.Eratio                                      */
    double minEfrac;          /* uinfo.minEfrac                                    */
    int unbindtype;           /* uinfo.unbindtype: 0 USYSANDPART, 1 UPART          */
    int bgpot;                /* uinfo.bgpot                                       */
    int cmvelreftype;         /* uinfo.cmvelreftype                                */
    double fracpotref;        /* uinfo.fracpotref                                  */
    long long Npotref;        /* uinfo.Npotref                                     */
    double maxunbindfrac;     /* uinfo.maxunbindfrac                               */
    double maxunboundfracforiterativeunbind;
    double maxallowedunboundfrac;
    double TreeThetaOpen;     /* uinfo.TreeThetaOpen                               */
    int BucketSize;           /* uinfo.BucketSize                                  */
    int verbose;
    int iterate_cm;           /* Options::iIterateCM (allvars.h:714, default 1)    */
    int sort_by_binding_energy; /* Options::iSortByBindingEnergy (allvars.h:750, default 1) */
} vrref_opts;

static void fill_options(Options &opt, const vrref_opts *o)
{
    opt.p = o->period;
    opt.ellxscale = o->ellxscale;
    opt.ellphys = o->ellphys;
    opt.ellhalophysfac = o->ellhalophysfac;
    opt.ellhalo6dxfac = o->ellhalo6dxfac;
    opt.ellhalo6dvfac = o->ellhalo6dvfac;
    opt.HaloMinSize = o->halo_min_size;
    opt.MinSize = o->min_size;
    opt.fofbgtype = o->fofbgtype;
    opt.iBoundHalos = o->bound_halos;
    opt.iopenmpfof = o->omp_fof;
    if (o->omp_fof_size > 0) opt.openmpfofsize = o->omp_fof_size;
    opt.iSubSearch = 0;
    opt.iKeepFOF = 0;
    opt.iBaryonSearch = 0;
    opt.partsearchtype = PSTALL;
    opt.iverbose = o->verbose;
    opt.G = o->G;
    opt.uinfo.eps = o->eps;
    opt.uinfo.Eratio = o->Eratio;
    opt.uinfo.minEfrac = o->minEfrac;
    opt.uinfo.unbindtype = o->unbindtype;
    opt.uinfo.bgpot = o->bgpot;
    opt.uinfo.cmvelreftype = o->cmvelreftype;
    opt.uinfo.fracpotref = o->fracpotref;
    opt.uinfo.Npotref = o->Npotref;
    opt.uinfo.maxunbindfrac = o->maxunbindfrac;
    opt.uinfo.maxunboundfracforiterativeunbind = o->maxunboundfracforiterativeunbind;
    opt.uinfo.maxallowedunboundfrac = o->maxallowedunboundfrac;
    opt.uinfo.TreeThetaOpen = o->TreeThetaOpen;
    opt.uinfo.BucketSize = o->BucketSize;
    opt.uinfo.icalculatepotential = true;
    opt.uinfo.unbindflag = 1;
    opt.iIterateCM = o->iterate_cm;
    opt.iSortByBindingEnergy = o->sort_by_binding_energy;
}

static void fill_particles(vector<Particle> &P, long long n, const double *pos, const double *vel,
                           const double *mass, const long long *pid)
{
    P.resize(n);
    for (long long i = 0; i < n; i++) {
        P[i].SetPosition(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        if (vel) P[i].SetVelocity(vel[3 * i], vel[3 * i + 1], vel[3 * i + 2]);
        else P[i].SetVelocity(0., 0., 0.);
        P[i].SetMass(mass ? mass[i] : 1.0);
        P[i].SetPID(pid ? pid[i] : i);
        P[i].SetID(i);
        P[i].SetType(DARKTYPE);
        P[i].SetDensity(0.);
        P[i].SetPotential(0.);
    }
}

size_t vrref_sizeof_particle(void) { return sizeof(Particle); }
int vrref_has_openmp(void)
{
#ifdef USEOPENMP
    return 1;
#else
    return 0;
#endif
}
int vrref_unbindnum(void) { return UNBINDNUM; }

/* Direct call of KDTree::FOF.  ndim = 3 (TPHYS, optional period) or 6 (TPHS, non periodic).
 * Returns number of groups; pfof_out[i] for input index i (order=1 -> size ordered). */
long long vrref_kdfof(long long n, const double *pos, const double *vel, int ndim, double fdist,
                      const double *period /*NULL or [3]*/, long long minnum, int bucket_size,
                      long long *pfof_out)
{
    vector<Particle> P;
    fill_particles(P, n, pos, vel, NULL, NULL);
    Double_t per[3];
    Double_t *pp = NULL;
    if (period) { for (int k = 0; k < 3; k++) per[k] = period[k]; pp = per; }
    KDTree *tree;
    if (ndim == 6) tree = new KDTree(P.data(), n, bucket_size, tree->TPHS, tree->KEPAN, 100);
    else tree = new KDTree(P.data(), n, bucket_size, tree->TPHYS, tree->KEPAN, 100, 0, 0, 0, pp);
    Int_t ng = 0;
    Int_t *pfof = tree->FOF(fdist, ng, minnum, 1);
    delete tree; /* restores input order */
    for (long long i = 0; i < n; i++) pfof_out[i] = pfof[i];
    delete[] pfof;
    return ng;
}

/* Direct call of KDTree::FOFCriterion(FOF6d, params, ng, minsize, order = 1, ipcheckflag = 1, FOFcheckbg) on a physical
 * tree, exactly as the 6D core search of SearchSubset does (search.cxx:1607-1611): params[1] = params[6] = p6,
 * params[2] = params[7] = p7, params[9] = p9, params[0] = tree type; the particles' potential field carries the mask. */
long long vrref_fofcriterion6d(long long n, const double *pos, const double *vel, const double *pot, double p6, double p7,
                               double p9, long long minsize, int bucket_size, long long *pfof_out)
{
    vector<Particle> P;
    fill_particles(P, n, pos, vel, NULL, NULL);
    for (long long i = 0; i < n; i++) P[i].SetPotential(pot[i]);
    KDTree *tree = new KDTree(P.data(), n, bucket_size, tree->TPHYS, tree->KEPAN, 100);
    Double_t param[20];
    for (int k = 0; k < 20; k++) param[k] = 0;
    param[0] = tree->GetTreeType();
    param[1] = param[6] = p6;
    param[2] = param[7] = p7;
    param[9] = p9;
    Int_t ng = 0;
#ifdef VRREF_USE_GPU
    /* the drop-in body gets the particles as the tree left them (reordered, ids = input index) */
    Int_t *pfof = FOFCriterion6DCoreGPU(P.data(), n, param, ng, minsize);
#else
    Int_t *pfof = tree->FOFCriterion(FOF6d, param, ng, minsize, 1, 1, FOFcheckbg);
#endif
    delete tree; /* restores input order */
    for (long long i = 0; i < n; i++) pfof_out[i] = pfof[i];
    delete[] pfof;
    return ng;
}

/* SearchFullSet on flat arrays.  pos/vel are IN/OUT (the reference unwraps periodic positions and
 * perturbs coordinates by the 6DFOF scale/unscale).  Returns number of groups.
 * seconds_out (optional) receives the wall time of the SearchFullSet call itself. */
long long vrref_search_full_set(long long n, double *pos, double *vel, const double *mass, const long long *pid,
                                const vrref_opts *o, long long *pfof_out, double *seconds_out)
{
    Options opt;
    fill_options(opt, o);
    vector<Particle> P;
    fill_particles(P, n, pos, vel, mass, pid);
    Int_t ng = 0;
    streambuf *old = NULL;
    stringstream sink;
    if (!o->verbose) old = cout.rdbuf(sink.rdbuf());
    auto t0 = chrono::steady_clock::now();
    Int_t *pfof = SearchFullSet(opt, n, P, ng);
    auto t1 = chrono::steady_clock::now();
    if (old) cout.rdbuf(old);
    if (seconds_out) *seconds_out = chrono::duration<double>(t1 - t0).count();
    for (long long i = 0; i < n; i++) {
        pfof_out[i] = pfof[i];
        for (int k = 0; k < 3; k++) { pos[3 * i + k] = P[i].GetPosition(k); vel[3 * i + k] = P[i].GetVelocity(k); }
    }
    delete[] pfof;
    return ng;
}

/* GetVelocityDensity (localfield.cxx:18) with tree == NULL, as SearchSubSub calls it for the particles of one structure
 * (search.cxx:2646): the routine builds its own physical tree (bucket Options::Bsize, Epanechnikov kernel table of 1000
 * entries, period Options::p) and deletes it again, which restores the input order.  approx = Options::
 * iLocalVelDenApproxCalcFlag (1, the default: one neighbour search per leaf node; 0: one per particle); approx = 2 is the
 * per-leaf search with the tree built by the caller and handed in (search.cxx:232).  type (may be NULL)
 * sets Particle::type -- this build defines STRUCDEN, so particles of type 0 get no density.  den_out in input order. */
void vrref_velocity_density(long long n, const double *pos, const double *vel, const int *type, double period, int bsize,
                            int nsearch, int nvel, int approx, double *den_out)
{
    Options opt;
    opt.p = period;
    opt.Bsize = bsize;
    opt.Nsearch = nsearch;
    opt.Nvel = nvel;
    opt.iLocalVelDenApproxCalcFlag = approx;
    opt.iBaryonSearch = 0;
    opt.partsearchtype = PSTALL;
    opt.iverbose = 0;
    vector<Particle> P;
    fill_particles(P, n, pos, vel, NULL, NULL);
    if (type) for (long long i = 0; i < n; i++) P[i].SetType(type[i]);
    stringstream sink;
    streambuf *old = cout.rdbuf(sink.rdbuf());
    if (approx == 1) GetVelocityDensity(opt, n, P.data(), NULL);
    else {
        /* approx == 2: the per-leaf search on a tree the caller holds, as SearchFullSet calls it (search.cxx:232) */
        opt.iLocalVelDenApproxCalcFlag = approx ? 1 : 0;
        /* GetVelocityDensityExact never builds a tree itself (it dereferences the NULL pointer): hand it one, as
         * SearchFullSet does with its FOF tree (search.cxx:232), built the way the approximative routine would */
        Double_t per[3] = {period, period, period};
        KDTree *tree = new KDTree(P.data(), n, bsize, tree->TPHYS, tree->KEPAN, 1000, 0, 0, 0, period > 0 ? per : NULL);
        GetVelocityDensity(opt, n, P.data(), tree);
        delete tree;
    }
    cout.rdbuf(old);
    for (long long i = 0; i < n; i++) den_out[P[i].GetID()] = P[i].GetDensity();
}

/* Tree potential of one group (Potential(), unbind.cxx:1607).  phi_out in input order. */
void vrref_potential(long long n, const double *pos, const double *mass, const vrref_opts *o, double *phi_out)
{
    Options opt;
    fill_options(opt, o);
    vector<Particle> P;
    fill_particles(P, n, pos, NULL, mass, NULL);
    Potential(opt, n, P.data());
    for (long long i = 0; i < n; i++) phi_out[i] = P[i].GetPotential();
}

/* CheckUnboundGroups on flat arrays. pfof is IN/OUT, returns new number of groups; *changed = return flag. */
long long vrref_unbind(long long n, const double *pos, const double *vel, const double *mass, const vrref_opts *o,
                       long long *pfof, long long ngroup, int *changed, double *seconds_out)
{
    Options opt;
    fill_options(opt, o);
    vector<Particle> P;
    fill_particles(P, n, pos, vel, mass, NULL);
    Int_t *pf = new Int_t[n];
    for (long long i = 0; i < n; i++) pf[i] = pfof[i];
    Int_t ng = ngroup;
    streambuf *old = NULL;
    stringstream sink;
    if (!o->verbose) old = cout.rdbuf(sink.rdbuf());
    auto t0 = chrono::steady_clock::now();
    int flag = CheckUnboundGroups(opt, n, P.data(), ng, pf);
    auto t1 = chrono::steady_clock::now();
    if (old) cout.rdbuf(old);
    if (seconds_out) *seconds_out = chrono::duration<double>(t1 - t0).count();
    if (changed) *changed = flag;
    for (long long i = 0; i < n; i++) pfof[i] = pf[i];
    delete[] pf;
    return ng;
}

/* CheckUnboundGroups with the full argument list of the SearchSubSub call sites (search.cxx:2674, :2867): numingroup and pglist
 * built by the reference's own BuildNumInGroup / BuildPGList, ireorder as given, groupflag (may be NULL, ngroup + 1 entries, IN/OUT).
 * numingroup_out: ngroup + 1 entries as the callee leaves them. */
long long vrref_unbind_full(long long n, const double *pos, const double *vel, const double *mass, const vrref_opts *o, long long *pfof,
                            long long ngroup, int ireorder, long long *groupflag, long long *numingroup_out, int *changed)
{
    Options opt;
    fill_options(opt, o);
    vector<Particle> P;
    fill_particles(P, n, pos, vel, mass, NULL);
    Int_t *pf = new Int_t[n];
    for (long long i = 0; i < n; i++) pf[i] = pfof[i];
    Int_t ng = ngroup;
    Int_t *num = BuildNumInGroup(n, ng, pf);
    Int_t **pgl = BuildPGList(n, ng, num, pf);
    Int_t *gf = NULL;
    if (groupflag) { gf = new Int_t[ngroup + 1]; for (long long g = 0; g <= ngroup; g++) gf[g] = groupflag[g]; }
    streambuf *old = NULL;
    stringstream sink;
    if (!o->verbose) old = cout.rdbuf(sink.rdbuf());
    int flag = CheckUnboundGroups(opt, n, P.data(), ng, pf, num, pgl, ireorder, gf);
    if (old) cout.rdbuf(old);
    if (changed) *changed = flag;
    for (long long i = 0; i < n; i++) pfof[i] = pf[i];
    for (long long g = 0; g <= ngroup; g++) { if (numingroup_out) numingroup_out[g] = num[g]; if (gf) groupflag[g] = gf[g]; }
    for (long long g = 1; g <= ngroup; g++) delete[] pgl[g];
    delete[] pgl; delete[] num; delete[] pf; delete[] gf;
    return ng;
}

/* GetBindingEnergy on GROUP-ORDERED flat arrays: group g = 1..ngroup owns numingroup[g] consecutive particles (the layout
 * SortAccordingtoBindingEnergy builds, substructureproperties.cxx:4279-4290).  gcmvel[3*g..] is stored into pdata[g].gcmvel
 * (what GetCM would have left there).  Outputs per particle: potential and binding energy (Particle::rho); per group: T,
 * Pot, Efrac and the positions inside the group of the minimum-potential and most-bound particles. */
void vrref_binding_energy(long long n, const double *pos, const double *vel, const double *mass, long long ngroup,
                          const long long *numingroup_in /* ngroup + 1 */, const double *gcmvel, const vrref_opts *o, double *phi_out,
                          double *E_out, double *T_out, double *Pot_out, double *Efrac_out, long long *iminpot_out, long long *ibound_out)
{
    Options opt;
    fill_options(opt, o);
    opt.ParticleTypeForRefenceFrame = -1;
    vector<Particle> P;
    fill_particles(P, n, pos, vel, mass, NULL); /* pid = id = position in the group-ordered array */
    Int_t *pfof = new Int_t[n];
    Int_t *numingroup = new Int_t[ngroup + 1];
    Int_t *noffset = new Int_t[ngroup + 1];
    numingroup[0] = 0;
    for (long long g = 1; g <= ngroup; g++) numingroup[g] = numingroup_in[g];
    if (ngroup >= 1) noffset[0] = noffset[1] = 0;
    for (long long g = 2; g <= ngroup; g++) noffset[g] = noffset[g - 1] + numingroup[g - 1];
    for (long long g = 1; g <= ngroup; g++)
        for (Int_t j = 0; j < numingroup[g]; j++) pfof[noffset[g] + j] = g;
    PropData *pdata = new PropData[ngroup + 1];
    for (long long g = 1; g <= ngroup; g++) {
        pdata[g].num = numingroup[g];
        for (int k = 0; k < 3; k++) pdata[g].gcmvel[k] = gcmvel[3 * g + k];
        pdata[g].T = pdata[g].Pot = pdata[g].Efrac = 0;
    }
    streambuf *old = NULL;
    stringstream sink;
    if (!o->verbose) old = cout.rdbuf(sink.rdbuf());
    GetBindingEnergy(opt, n, P.data(), ngroup, pfof, numingroup, pdata, noffset);
    if (old) cout.rdbuf(old);
    for (long long i = 0; i < n; i++) { phi_out[i] = P[i].GetPotential(); E_out[i] = P[i].GetDensity(); }
    for (long long g = 1; g <= ngroup; g++) {
        T_out[g] = pdata[g].T; Pot_out[g] = pdata[g].Pot; Efrac_out[g] = pdata[g].Efrac;
        iminpot_out[g] = pdata[g].iminpot - noffset[g];
        ibound_out[g] = pdata[g].ibound - noffset[g];
    }
    delete[] pdata;
    delete[] pfof;
    delete[] numingroup;
    delete[] noffset;
}

/* SortAccordingtoBindingEnergy (src/substructureproperties.cxx:4256-4352): pfof by particle index in, per group the PIDs
 * and total energies in the order the reference leaves the particles in, numingroup and pglist[g][numingroup[g]]
 * (= pdata[g].iunbound) out.  libvrref.so runs the reference's function whole.  libvrref_gpu.so runs the flow a
 * maintainer gets after the INTEGRATION.md patch: particles brought into group order, the reference's own GetCM, then
 * GetBindingEnergy and the per-group sort on the B200 (the other property calculations of the reference do not feed the
 * ordering and are left out of this harness). */
void vrref_sort_by_binding_energy(long long n, const double *pos, const double *vel, const double *mass, const long long *pid,
                                  const long long *pfof_in, long long ngroup, const vrref_opts *o, long long *pid_sorted,
                                  double *E_sorted, long long *numingroup_out /* ngroup + 1 */, long long *iunbound_out /* ngroup + 1 */)
{
    Options opt;
    fill_options(opt, o);
    vector<Particle> P;
    fill_particles(P, n, pos, vel, mass, pid);
    Int_t *pfof = new Int_t[n > 0 ? n : 1];
    Int_t *numingroup = new Int_t[ngroup + 1];
    for (long long g = 0; g <= ngroup; g++) numingroup[g] = 0;
    for (long long i = 0; i < n; i++) { pfof[i] = pfof_in[i]; if (pfof[i] > 0) numingroup[pfof[i]]++; }
    PropData *pdata = new PropData[ngroup + 1];
    /* every group is a field halo: what CopyHierarchy (search.cxx:3658-3659) records for level-0 objects */
    for (long long g = 0; g <= ngroup; g++) pdata[g].hostid = pdata[g].directhostid = GROUPNOPARENT;
    streambuf *old = NULL;
    stringstream sink;
    if (!o->verbose) old = cout.rdbuf(sink.rdbuf());
    long long k = 0;
#ifdef VRREF_USE_GPU
    Int_t *noffset = new Int_t[ngroup + 2];
    noffset[0] = 0;
    if (ngroup >= 1) noffset[1] = 0;
    for (long long g = 2; g <= ngroup + 1; g++) noffset[g] = noffset[g - 1] + numingroup[g - 1];
    {
        /* group order: grouped particles by ascending group id, members by ascending index, the rest behind them */
        vector<Particle> Q(n);
        vector<Int_t> fill(noffset, noffset + ngroup + 2);
        Int_t rest = ngroup >= 1 ? noffset[ngroup + 1] : 0;
        for (long long i = 0; i < n; i++) Q[pfof[i] > 0 ? fill[pfof[i]]++ : rest++] = P[i];
        P.swap(Q);
    }
    for (long long g = 1; g <= ngroup; g++) pdata[g].num = numingroup[g];
    GetCM(opt, n, P.data(), ngroup, pfof, numingroup, pdata, noffset);
    GetBindingEnergy(opt, n, P.data(), ngroup, pfof, numingroup, pdata, noffset);
    SortGroupParticlesGPU(opt, P.data(), ngroup, numingroup, noffset, pdata);
    for (long long g = 1; g <= ngroup; g++) {
        for (Int_t j = 0; j < numingroup[g]; j++, k++) {
            pid_sorted[k] = P[noffset[g] + j].GetPID();
            E_sorted[k] = P[noffset[g] + j].GetDensity();
        }
        iunbound_out[g] = pdata[g].iunbound;
    }
    delete[] noffset;
#else
    Int_t **pglist = SortAccordingtoBindingEnergy(opt, n, P.data(), ngroup, pfof, numingroup, pdata);
    for (long long g = 1; g <= ngroup; g++) {
        for (Int_t j = 0; j < numingroup[g]; j++, k++) {
            pid_sorted[k] = P[pglist[g][j]].GetPID();
            E_sorted[k] = P[pglist[g][j]].GetDensity();
        }
        iunbound_out[g] = pglist[g][numingroup[g] > 0 ? numingroup[g] : 0];
        delete[] pglist[g];
    }
    delete[] pglist;
#endif
    if (old) cout.rdbuf(old);
    numingroup_out[0] = 0;
    for (long long g = 1; g <= ngroup; g++) numingroup_out[g] = numingroup[g];
    delete[] pdata;
    delete[] numingroup;
    delete[] pfof;
}

/* BuildNumInGroup + BuildPGList (src/buildandsortarrays.cxx:11-16, 26-43): sizes[0..ngroups] and the lists of group 1, 2, ...
 * written one after the other into index_out (n entries available); returns the number of grouped particles. */
long long vrref_build_pglist(long long n, const long long *pfof_in, long long ngroups, long long *numingroup_out, long long *index_out)
{
    Int_t *pfof = new Int_t[n > 0 ? n : 1];
    for (long long i = 0; i < n; i++) pfof[i] = pfof_in[i];
    Int_t *numingroup = BuildNumInGroup(n, ngroups, pfof);
    Int_t **pglist = BuildPGList(n, ngroups, numingroup, pfof);
    long long k = 0;
    for (long long g = 0; g <= ngroups; g++) numingroup_out[g] = numingroup[g];
    for (long long g = 1; g <= ngroups; g++) {
        for (Int_t j = 0; j < numingroup[g]; j++) index_out[k++] = pglist[g][j];
        if (pglist[g] != NULL) delete[] pglist[g];
    }
    delete[] pglist;
    delete[] numingroup;
    delete[] pfof;
    return k;
}

} /* extern "C" */
