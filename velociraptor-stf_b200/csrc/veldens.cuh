// velociraptor-stf_b200/csrc/veldens.cuh -- local velocity density, per-thread code.
//
// GetVelocityDensity (src/localfield.cxx:18) for the particles of one structure, as SearchSubSub calls it (tree == NULL):
//   approx = 1  GetVelocityDensityApproximative (:726-897, the default): per LEAF of the structure's kd-tree (bucket
//               Options::Bsize) the centre of its typed particles, the Nsearch nearest particles of that centre
//               (KDTree::FindNearestPos), and for every typed particle of the leaf the Nvel nearest of those in velocity,
//               itself excluded;
//   approx = 0  GetVelocityDensityExact (:485-586): the Nsearch nearest particles of every typed particle
//               (KDTree::FindNearest(target): the target and, in an open domain, every particle at distance zero left out);
// then KDTree::CalcSmoothLocalValue (NBodylib/src/KDTree/KDCalcSmoothQuantities.cxx:1704-1719): the Epanechnikov kernel
// table interpolated at the Nvel velocity distances, summed from the largest distance down.
//
// The leaves are the reference's (the tree comes from the exact-ties build, qsel.cuh); the neighbour search returns a SET
// -- the Nsearch smallest squared distances -- so it walks the same tree in its own order: depth first with a small
// stack, nearer child first, a subtree skipped when its bounding box is farther than the current Nsearch-th distance
// (rounding is monotone, so a box distance never exceeds the distance of a particle inside it).  Periodic domains run
// the reference's sequence of up to eight reflected query points into the same queue with its gates.
//
// First version: ONE THREAD per query with its queue in global scratch, so that the same source runs as a serial g++
// emulation against the oracle (tests/host/veldens_emul.cpp); the warp-cooperative search is the next step once this
// one has been seen on the GPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define VD_HD __device__ __forceinline__
#define VD_MEMBER __device__ __forceinline__
#else
#include <math.h>
#define VD_HD static inline
#define VD_MEMBER inline
#endif

namespace vrgpu {
namespace veldens {

constexpr uint32_t NONE = 0xffffffffu;
constexpr double MAXVALUE = 1e32;   // NBodylib/src/Math/Precision.h:39
constexpr int MAXVEL = 64;          // largest Nvel (the queue of the velocity step lives in registers / local memory)
constexpr int STACK = 96;           // depth of the search stack (a median-split tree of 2^32 particles is 32 deep)

#if defined(__CUDACC__)
VD_HD double fsub(double a, double b) { return __dsub_rn(a, b); }
VD_HD double fadd(double a, double b) { return __dadd_rn(a, b); }
VD_HD double fmul(double a, double b) { return __dmul_rn(a, b); }
VD_HD double fdiv(double a, double b) { return __ddiv_rn(a, b); }
VD_HD double fsqrt(double a) { return __dsqrt_rn(a); }
// hi^3 with one rounding (the reference calls pow(hi, 3.0), which glibc rounds correctly up to ~2^-13 of the cases)
VD_HD double cube(double h)
{
    const double p = __dmul_rn(h, h), e = __fma_rn(h, h, -p);
    return __fma_rn(p, h, __dmul_rn(e, h));
}
#else
VD_HD double fsub(double a, double b) { return a - b; }
VD_HD double fadd(double a, double b) { return a + b; }
VD_HD double fmul(double a, double b) { return a * b; }
VD_HD double fdiv(double a, double b) { return a / b; }
VD_HD double fsqrt(double a) { return sqrt(a); }
VD_HD double cube(double h)
{
    const double p = h * h, e = fma(h, h, -p);
    return fma(p, h, e * h);
}
#endif

struct Params {
    // the structure's particles in tree order
    const double *tx, *ty, *tz, *tvx, *tvy, *tvz;
    const int *ttype;               // Particle::type (0: no density, STRUCDEN)
    uint32_t n;
    // tree (ptree.cuh numbering); bb = 6 doubles per node: lower corner, upper corner
    const uint32_t *nstart, *nend, *nleft;
    const double *bb;
    uint32_t root;
    // leaves
    const uint32_t *leaf;           // leaf index -> node
    uint32_t *leaf_of;              // tree position -> leaf index
    double *cm;                     // 3 per leaf
    uint32_t *cnum;                 // typed particles per leaf
    uint32_t nleaf;
    // search
    double period;                  // 0: open domain
    uint32_t nsearch, nvel;
    int approx;
    // one batch of queries: [q0, q1) are leaf indices (approx) or tree positions (exact)
    uint32_t q0, q1;
    double *heap_d2;                // (nsearch + 1) * (q1 - q0): entry i of slot s at [i * (q1 - q0) + s]
    uint32_t *heap_id;
    uint32_t *nn;                   // nsearch * (q1 - q0), same layout: the neighbours, NONE where there are fewer
    // kernel table (KDTree::KernelConstruction, KDTree.cxx:592-631)
    const double *kernel;
    int kernres;
    double *den;                    // per tree position
    unsigned int *overflow;         // set when a search would have needed more than STACK entries (null: not reported)
};

// ---- leaves ------------------------------------------------------------------------------------------------------------
// localfield.cxx:791-806: centre of the typed particles, sums in tree order, divided by (float)num
VD_HD void leaf_thread(const Params &P, uint32_t l)
{
    const uint32_t nd = P.leaf[l];
    uint32_t num = 0;
    double cx = 0, cy = 0, cz = 0;
    for (uint32_t j = P.nstart[nd]; j < P.nend[nd]; j++) {
        P.leaf_of[j] = l;
        if (P.ttype[j] == 0) continue;
        num++;
        cx = fadd(cx, P.tx[j]); cy = fadd(cy, P.ty[j]); cz = fadd(cz, P.tz[j]);
    }
    P.cnum[l] = num;
    if (num) {
        const double d = (double)(float)num;
        cx = fdiv(cx, d); cy = fdiv(cy, d); cz = fdiv(cz, d);
    }
    P.cm[3 * l] = cx; P.cm[3 * l + 1] = cy; P.cm[3 * l + 2] = cz;
}

// Leaves handed in by the caller (the reference passes its FOF tree over the same particles, search.cxx:232: the particles
// are then in that tree's order and leaf l is the index range [off[l], off[l+1])): centre and typed count from the caller's
// array in the caller's order, exactly the sums of localfield.cxx:791-806 on that tree.
struct CallerLeaves {
    const double *pos;              // packed xyz in the caller's order
    const int *type;                // caller's order, may be null (all typed)
    const long long *off;           // nleaf + 1
    uint32_t nleaf;
    uint32_t *leaf_of_orig;         // caller index -> leaf
    double *cm;
    uint32_t *cnum;
};
VD_HD void caller_leaf_thread(const CallerLeaves &C, uint32_t l)
{
    uint32_t num = 0;
    double cx = 0, cy = 0, cz = 0;
    for (long long i = C.off[l]; i < C.off[l + 1]; i++) {
        C.leaf_of_orig[i] = l;
        if (C.type && C.type[i] == 0) continue;
        num++;
        cx = fadd(cx, C.pos[3 * i]); cy = fadd(cy, C.pos[3 * i + 1]); cz = fadd(cz, C.pos[3 * i + 2]);
    }
    C.cnum[l] = num;
    if (num) {
        const double d = (double)(float)num;
        cx = fdiv(cx, d); cy = fdiv(cy, d); cz = fdiv(cz, d);
    }
    C.cm[3 * l] = cx; C.cm[3 * l + 1] = cy; C.cm[3 * l + 2] = cz;
}

// ---- the queue: a binary max-heap of K (distance, id) pairs --------------------------------------------------------------
// Entry i of the queue of query slot s lives at [i * stride + s] (stride = queries of the batch): the threads of a warp own
// consecutive slots, so their accesses to the same heap position -- the top above all -- fall into consecutive words.
struct Heap {
    double *d2;
    uint32_t *id;
    uint32_t K;
    size_t stride;
    VD_MEMBER double &D(uint32_t i) const { return d2[i * stride]; }
    VD_MEMBER uint32_t &I(uint32_t i) const { return id[i * stride]; }
};
VD_HD void heap_init(const Heap &H)
{
    for (uint32_t i = 0; i < H.K; i++) { H.D(i) = MAXVALUE; H.I(i) = NONE; }
}
// Entries are ordered by (distance, id): of several particles at exactly the Nsearch-th distance (float32 coordinates
// produce such ties) the ones with the lower tree position are kept, whatever the order of the walk -- the oracle's rule.
VD_HD bool after(double d2a, uint32_t ida, double d2b, uint32_t idb) { return d2a > d2b || (d2a == d2b && ida > idb); }
VD_HD void heap_replace_top(const Heap &H, double d2, uint32_t id)
{
    uint32_t i = 0;
    for (;;) {
        const uint32_t l = 2 * i + 1, r = l + 1;
        if (l >= H.K) break;
        const uint32_t c = (r < H.K && after(H.D(r), H.I(r), H.D(l), H.I(l))) ? r : l;
        if (!after(H.D(c), H.I(c), d2, id)) break;
        H.D(i) = H.D(c); H.I(i) = H.I(c);
        i = c;
    }
    H.D(i) = d2; H.I(i) = id;
}

// squared distance from q to the bounding box of node nd, in the operation order of a particle distance
VD_HD double box_d2(const Params &P, uint32_t nd, const double *q)
{
    const double *b = P.bb + 6 * (size_t)nd;
    double s = 0;
    for (int k = 0; k < 3; k++) {
        double d = 0;
        if (q[k] < b[k]) d = fsub(q[k], b[k]);
        else if (q[k] > b[3 + k]) d = fsub(q[k], b[3 + k]);
        s = fadd(s, fmul(d, d));
    }
    return s;
}

// one tree search from the point q: what KDTree::FindNearestPos / FindNearest does to the priority queue.
// target != NONE (open domain only): FindNearest(target) -- the target and particles at distance zero are not offered.
VD_HD void search_from(const Params &P, const Heap &H, const double *q, uint32_t target)
{
    uint32_t stack[STACK];
    int sp = 0;
    stack[sp++] = P.root;
    while (sp > 0) {
        const uint32_t nd = stack[--sp];
        if (box_d2(P, nd, q) > H.D(0)) continue;
        const uint32_t l = P.nleft[nd];
        if (l == NONE) {
            for (uint32_t i = P.nstart[nd]; i < P.nend[nd]; i++) {
                if (i == target) continue;
                const double dx = fsub(q[0], P.tx[i]), dy = fsub(q[1], P.ty[i]), dz = fsub(q[2], P.tz[i]);
                const double d2 = fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz));
                if (target != NONE && !(d2 > 0)) continue;
                if (after(H.D(0), H.I(0), d2, i)) heap_replace_top(H, d2, i);
            }
            continue;
        }
        const double dl = box_d2(P, l, q), dr = box_d2(P, l + 1u, q);
        // the nearer child is searched first: push the farther one first
        if (sp >= STACK - 1) { // cannot happen for a median-split tree (32 levels): reported, never silently dropped
            if (P.overflow) *P.overflow = 1u;
            continue;
        }
        if (dl <= dr) { stack[sp++] = l + 1u; stack[sp++] = l; }
        else { stack[sp++] = l; stack[sp++] = l + 1u; }
    }
}

// the neighbours of one query: leaf centre (approx) or particle (exact)
VD_HD void knn_thread(const Params &P, uint32_t q)
{
    const uint32_t slot = q - P.q0;
    const size_t stride = P.q1 - P.q0;
    uint32_t *out = P.nn + slot;                                   // neighbour k at out[k * stride]
    double c[3];
    uint32_t target = NONE;
    if (P.approx) {
        if (P.cnum[q] == 0) return;
        c[0] = P.cm[3 * q]; c[1] = P.cm[3 * q + 1]; c[2] = P.cm[3 * q + 2];
    } else {
        if (P.ttype[q] == 0) return;
        c[0] = P.tx[q]; c[1] = P.ty[q]; c[2] = P.tz[q];
        target = q;
    }
    const bool periodic = P.period > 0;
    const uint32_t K = P.nsearch + ((periodic && target != NONE) ? 1u : 0u);
    Heap H{P.heap_d2 + slot, P.heap_id + slot, K, stride};
    heap_init(H);
    if (!periodic) {
        search_from(P, H, c, target);
        for (uint32_t k = 0; k < P.nsearch; k++) out[k * stride] = H.I(k);
        return;
    }
    // SplitNode::FindNearestPosPeriodic (KDSplitNode.cxx:1008-1040): the point, then its reflections, each behind its gate
    search_from(P, H, c, NONE);
    double xp[3], sval;
    const double half = fdiv(P.period, 2.0);
    for (int k = 0; k < 3; k++) {                                  // PeriodicReflection1D (DistFunc.h:326-332)
        xp[0] = c[0]; xp[1] = c[1]; xp[2] = c[2];
        if (c[k] < half) { xp[k] = fadd(c[k], P.period); sval = c[k]; }
        else { xp[k] = fsub(c[k], P.period); sval = -xp[k]; }
        if (fsqrt(H.D(0)) > sval) search_from(P, H, xp, NONE);
    }
    for (int r = 0; r <= 3; r++) {                                 // PeriodicReflection2D (0,1) (0,2) (1,2), then ND
        xp[0] = c[0]; xp[1] = c[1]; xp[2] = c[2];
        sval = 0;
        const int na = r < 3 ? 2 : 3;
        for (int a = 0; a < na; a++) {
            const int k = r == 0 ? a : (r == 1 ? 2 * a : (r == 2 ? a + 1 : a));
            if (c[k] < half) { xp[k] = fadd(c[k], P.period); sval = fadd(sval, fmul(c[k], c[k])); }
            else { xp[k] = fsub(c[k], P.period); sval = fadd(sval, fmul(xp[k], xp[k])); }
        }
        sval = fsqrt(sval);
        if (H.D(0) > sval) search_from(P, H, xp, NONE);            // the squared radius against a distance: Q16
    }
    if (target == NONE) {
        for (uint32_t k = 0; k < P.nsearch; k++) out[k * stride] = H.I(k);
        return;
    }
    // KDFindNearest.cxx:292, 356-357: the queue holds Nsearch + 1 entries and the smallest is not loaded
    uint32_t imin = 0;
    for (uint32_t k = 1; k < K; k++) if (after(H.D(imin), H.I(imin), H.D(k), H.I(k))) imin = k;
    uint32_t m = 0;
    for (uint32_t k = 0; k < K; k++) if (k != imin) out[(m++) * stride] = H.I(k);
}

// ---- velocity step + kernel sum, one typed particle ------------------------------------------------------------------------
VD_HD void density_thread(const Params &P, uint32_t j)
{
    const uint32_t q = P.approx ? P.leaf_of[j] : j;
    if (q < P.q0 || q >= P.q1) return;
    P.den[j] = 0.0;
    if (P.ttype[j] == 0) return;
    const uint32_t *ids = P.nn + (q - P.q0);
    const size_t stride = P.q1 - P.q0;
    double h[MAXVEL];
    const uint32_t nv = P.nvel;
    for (uint32_t k = 0; k < nv; k++) h[k] = MAXVALUE;
    uint32_t m = 0;
    const double vx = P.tvx[j], vy = P.tvy[j], vz = P.tvz[j];
    for (uint32_t k = 0; k < P.nsearch; k++) {
        const uint32_t id = ids[k * stride];
        if (id == NONE || (P.approx && id == j)) continue;
        const double ax = fsub(vx, P.tvx[id]), ay = fsub(vy, P.tvy[id]), az = fsub(vz, P.tvz[id]);
        const double v2 = fadd(fadd(fmul(ax, ax), fmul(ay, ay)), fmul(az, az));
        m++;
        if (!(v2 < h[0])) continue;
        uint32_t i = 0;                                            // replace the top of the max-heap
        for (;;) {
            const uint32_t l = 2 * i + 1, r = l + 1;
            if (l >= nv) break;
            const uint32_t c = (r < nv && h[r] > h[l]) ? r : l;
            if (!(h[c] > v2)) break;
            h[i] = h[c];
            i = c;
        }
        h[i] = v2;
    }
    if (m < nv) return;                                            // fewer neighbours than Nvel: no density (as the oracle)
    // CalcSmoothLocalValue: largest distance first
    const double hi = fmul(0.5, fsqrt(h[0]));
    const double norm = fdiv(1.0, cube(hi));
    const double delta = fdiv(2.0, (double)(P.kernres - 1));
    double value = 0;
    for (uint32_t left = nv; left > 0; left--) {
        const double v2 = h[0];
        // pop: move the last entry to the top and sift it down
        const double last = h[left - 1];
        uint32_t i = 0;
        for (;;) {
            const uint32_t l = 2 * i + 1, r = l + 1;
            if (l >= left - 1) break;
            const uint32_t c = (r < left - 1 && h[r] > h[l]) ? r : l;
            if (!(h[c] > last)) break;
            h[i] = h[c];
            i = c;
        }
        h[i] = last;
        const double rij = fsqrt(v2);
        const double r = fdiv(rij, hi);
        const int ik = (int)fmul(fmul(r, 0.5), (double)(P.kernres - 1));
        double w;
        if (ik < P.kernres - 1)
            w = fadd(P.kernel[ik], fdiv(fmul(fsub(P.kernel[ik + 1], P.kernel[ik]), fsub(r, fmul(delta, (double)ik))), delta));
        else w = P.kernel[ik];
        value = fadd(value, fmul(fmul(w, norm), 1.0));
    }
    P.den[j] = value;
}

} // namespace veldens
} // namespace vrgpu
