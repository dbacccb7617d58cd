// velociraptor-stf_b200/csrc/qsel.cuh -- the reference's median selection, arrangement included (VRGPU_TREE_TIES=1, ptree.cu).
//
// NBody::KDTree::BuildNodes (NBodylib/src/KDTree/KDTree.cxx:438-507) splits a node with MedianPos (:304-337), a quickselect
// with a Hoare partition about the element at the median index.  Which of several particles with EXACTLY equal split
// coordinate end up left of the median -- and in which order every particle is left for the next level -- follows from its
// swap sequence; float32-quantised inputs have such ties at most levels of a big group, and node membership decides which
// cells the walk opens.  The default device build sorts (ties -> lower member index: DESIGN.md "tree parity"); this
// variant reproduces the reference's arrangement itself, so node membership AND particle order are the reference's for
// every input.
//
// One iteration of MedianPos on the positions [lo, hi) of a node (pivot index k, pivot value x) is data parallel:
//     swap p[k] <-> p[hi-1]
//     A = positions of [lo, hi-1) holding a value >= x, increasing   (where the upward scan i stops)
//     B = positions of [lo, hi-1) holding a value <= x, decreasing   (where the downward scan j stops)
//     M = number of m with A[m] < B[m]   (a prefix of the pairs: A increases, B decreases);   swap p[A[m]] <-> p[B[m]], m < M
//     i* = A[M] if it exists and (M == 0 or A[M] < B[M-1]), else B[M-1] if M > 0, else hi-1;    swap p[i*] <-> p[hi-1]
//     hi = i* if i* >= k;  lo = i* + 1 if i* <= k;  repeat while lo + 1 < hi
// (the swapped positions are pairwise distinct: a position with value == x is in A and in B, but A[m] == B[m'] with
// m, m' < M contradicts A[m'] < B[m'] or A[m] < B[m]).  tests/test_hoare_model.py pins this form against a literal
// transcription of MedianPos; tests/host/qsel_emul.cpp runs the per-thread functions below (the same source the kernels
// of ptree.cu call) level by level and compares the whole arrangement with the oracle's tree.
//
// All splitting nodes of a level above QS_SMALL particles advance together, one "round" = one iteration of each.  The work
// items of a round are the scanned positions of the selections that still run, laid back to back (an exclusive scan of the
// remaining lengths gives every selection its first item and the host the launch size), so a round costs what MedianPos
// itself touches -- about 3.4 n per node over all rounds -- and not rounds x n.  Five launches and two scans per round;
// the ranks inside A and B come from one exclusive scan of the packed 64-bit flags.  Nodes of at most QS_SMALL particles
// are selected by one thread each on a copy in shared memory (select_serial: the literal loop).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define QS_HD __device__ __forceinline__
#else
#define QS_HD static inline
#endif

namespace vrgpu {
namespace qsel {

constexpr uint32_t QS_SMALL = 2048; // nodes up to this size: serial selection in shared memory (24 KB per CTA)

struct Level {
    const double *x, *y, *z;      // group-major columns
    const uint32_t *slot_of;      // compact member index -> slot in the columns
    uint32_t *perm;               // tree position -> compact member index (rearranged in place)
    const uint32_t *frontier;     // frontier index -> node
    const uint32_t *nstart, *nend;
    const uint32_t *splitflag, *splitdim;
    uint32_t *lo, *hi, *npair;    // per frontier entry: remaining range [lo, hi) of the selection, M of the round
    uint32_t *len;                // per frontier entry: positions scanned in the coming round (hi - lo - 1, 0 when finished)
    const uint32_t *woff;         // exclusive scan of len: the entry's first work item
    double *pivot;                // per frontier entry: x of the round
    // per work item w (the work items of a round are the scanned positions of all running selections, back to back)
    uint32_t *wf;                 // frontier entry of the item
    unsigned long long *flags;    // (value >= x) | (value <= x) << 32; W + 1 entries, the last one zero
    const unsigned long long *scan; // exclusive scan of flags (W + 1 entries: scan[W] = totals)
    uint32_t *A, *B;              // per entry, stored from work item woff[f]
    uint32_t nF, W;
};

QS_HD double coord(const Level &L, uint32_t d, uint32_t member)
{
    const uint32_t s = L.slot_of[member];
    return d == 0 ? L.x[s] : (d == 1 ? L.y[s] : L.z[s]);
}
QS_HD uint32_t median_index(uint32_t start, uint32_t end) { return start + (end - start - 1u) / 2u; }
QS_HD bool running(uint32_t lo, uint32_t hi) { return lo + 1u < hi; } // MedianPos: while (left < right)
QS_HD uint32_t lo32(unsigned long long v) { return (uint32_t)(v & 0xffffffffull); }
QS_HD uint32_t hi32(unsigned long long v) { return (uint32_t)(v >> 32); }

// the last entry f with woff[f] <= w: entries without work share their offset with the next one, so this is the entry
// that owns work item w (w < W, woff[0] == 0)
QS_HD uint32_t find_entry(const uint32_t *woff, uint32_t nF, uint32_t w)
{
    uint32_t a = 0u, b = nF;
    while (b - a > 1u) {
        const uint32_t m = a + (b - a) / 2u;
        if (woff[m] <= w) a = m; else b = m;
    }
    return a;
}

// ---- per frontier entry: start of a level ------------------------------------------------------------------------
QS_HD void init_thread(const Level &L, uint32_t f)
{
    const uint32_t nd = L.frontier[f];
    const uint32_t s = L.nstart[nd], e = L.nend[nd];
    const bool big = L.splitflag[f] && e - s > QS_SMALL;
    L.lo[f] = big ? s : 0u;
    L.hi[f] = big ? e : 0u;
    L.len[f] = big && running(s, e) ? e - s - 1u : 0u;
}

// ---- per frontier entry: pivot of the round ----------------------------------------------------------------------
QS_HD void begin_thread(const Level &L, uint32_t f)
{
    const uint32_t lo = L.lo[f], hi = L.hi[f];
    if (!running(lo, hi)) return;
    const uint32_t nd = L.frontier[f];
    const uint32_t k = median_index(L.nstart[nd], L.nend[nd]);
    const uint32_t pk = L.perm[k], pr = L.perm[hi - 1u];
    L.pivot[f] = coord(L, L.splitdim[f], pk);
    L.perm[k] = pr;
    L.perm[hi - 1u] = pk;
    L.npair[f] = 0u;
}

// ---- per work item (W + 1 threads): where the two scans stop -----------------------------------------------------
QS_HD void flag_thread(const Level &L, uint32_t w)
{
    if (w >= L.W) { L.flags[w] = 0ull; return; }
    const uint32_t f = find_entry(L.woff, L.nF, w);
    const uint32_t p = L.lo[f] + (w - L.woff[f]);
    const double v = coord(L, L.splitdim[f], L.perm[p]), x = L.pivot[f];
    L.flags[w] = (v >= x ? 1ull : 0ull) | (v <= x ? (1ull << 32) : 0ull);
    L.wf[w] = f;
}

// ---- per work item: A and B --------------------------------------------------------------------------------------
QS_HD void scatter_thread(const Level &L, uint32_t w)
{
    const unsigned long long fl = L.flags[w];
    if (fl == 0ull) return;
    const uint32_t f = L.wf[w];
    const uint32_t base = L.woff[f];
    const uint32_t p = L.lo[f] + (w - base);
    const unsigned long long s0 = L.scan[base], s1 = L.scan[base + L.len[f]], sw = L.scan[w];
    if (lo32(fl)) L.A[base + (lo32(sw) - lo32(s0))] = p;
    if (hi32(fl)) L.B[base + ((hi32(s1) - hi32(s0)) - 1u - (hi32(sw) - hi32(s0)))] = p;
}

// ---- per work item, read as pair number m = w - woff[f]: the swaps -----------------------------------------------
QS_HD void match_thread(const Level &L, uint32_t w)
{
    const uint32_t f = L.wf[w];
    const uint32_t base = L.woff[f];
    const unsigned long long s0 = L.scan[base], s1 = L.scan[base + L.len[f]];
    const uint32_t nA = lo32(s1) - lo32(s0), nB = hi32(s1) - hi32(s0);
    const uint32_t mm = nA < nB ? nA : nB, m = w - base;
    if (m >= mm) return;
    const uint32_t a = L.A[w], b = L.B[w];
    if (!(a < b)) return;
    const uint32_t pa = L.perm[a], pb = L.perm[b];
    L.perm[a] = pb;
    L.perm[b] = pa;
    if (m + 1u == mm || !(L.A[w + 1u] < L.B[w + 1u])) L.npair[f] = m + 1u;
}

// ---- per frontier entry: the pivot goes to its place, the range shrinks ------------------------------------------
QS_HD void end_thread(const Level &L, uint32_t f)
{
    const uint32_t lo = L.lo[f], hi = L.hi[f];
    if (!running(lo, hi)) return;
    const uint32_t nd = L.frontier[f];
    const uint32_t k = median_index(L.nstart[nd], L.nend[nd]);
    const uint32_t base = L.woff[f];
    const unsigned long long s0 = L.scan[base], s1 = L.scan[base + L.len[f]];
    const uint32_t nA = lo32(s1) - lo32(s0);
    const uint32_t M = L.npair[f];
    uint32_t istar;
    if (M < nA && (M == 0u || L.A[base + M] < L.B[base + M - 1u])) istar = L.A[base + M];
    else if (M > 0u) istar = L.B[base + M - 1u];
    else istar = hi - 1u;
    const uint32_t pv = L.perm[hi - 1u], pi = L.perm[istar];
    L.perm[hi - 1u] = pi;
    L.perm[istar] = pv;
    uint32_t nlo = lo, nhi = hi;
    if (istar >= k) nhi = istar;
    if (istar <= k) nlo = istar + 1u;
    L.lo[f] = nlo;
    L.hi[f] = nhi;
    L.len[f] = running(nlo, nhi) ? nhi - nlo - 1u : 0u;
}

// ---- small nodes: MedianPos itself on (value, member) pairs of one node, k relative to the node's start -----------
QS_HD void select_serial(double *val, uint32_t *idx, uint32_t k, uint32_t n)
{
    if (n == 0u) return;
    int left = 0, right = (int)n - 1;
    const int kk = (int)k;
    while (left < right) {
        const double x = val[kk];
        double wv = val[right]; val[right] = val[kk]; val[kk] = wv;
        uint32_t wi = idx[right]; idx[right] = idx[kk]; idx[kk] = wi;
        int i = left - 1, j = right;
        for (;;) {
            while (i < j) if (val[++i] >= x) break;
            while (i < j) if (val[--j] <= x) break;
            wv = val[i]; val[i] = val[j]; val[j] = wv;
            wi = idx[i]; idx[i] = idx[j]; idx[j] = wi;
            if (j <= i) break;
        }
        val[j] = val[i]; idx[j] = idx[i];
        val[i] = val[right]; idx[i] = idx[right];
        val[right] = wv; idx[right] = wi;
        if (i >= kk) right = i - 1;
        if (i <= kk) left = i + 1;
    }
}

// a frontier entry handled by select_serial on this level?
QS_HD bool small_node(const Level &L, uint32_t f, uint32_t &s, uint32_t &e)
{
    const uint32_t nd = L.frontier[f];
    s = L.nstart[nd];
    e = L.nend[nd];
    return L.splitflag[f] && e - s <= QS_SMALL;
}

} // namespace qsel
} // namespace vrgpu
