// velociraptor-stf_b200/csrc/ptree.cu -- kd-tree monopole ("Barnes-Hut") potentials for large groups.
//
// Replaces Potential() (src/unbind.cxx:1607-1741) with its helpers GetNodeList (:13-21) and MarkCell (:24-44) and the
// kd-tree it builds (NBody::KDTree::BuildNodes, NBodylib/src/KDTree/KDTree.cxx:438-507: split dimension = largest
// spread (first wins), median index k = start + (size-1)/2, left child = the k-start+1 smallest along that dimension,
// leaf when size <= BucketSize).
//
// Device design: every group above the direct-sum threshold is built in the SAME launches, level by level.
//   * ranks: three global radix sorts give every particle its rank along x, y, z (ties -> member order);
//   * a level = the "frontier" of current nodes, ordered by start.  Per level: segment index per particle (head flags +
//     scan), per-node min/max per axis (ordered-integer atomics), split dimension, then ONE stable radix sort on the
//     key (frontier index, rank along the node's split dimension) -- after it the first k-start+1 particles of each
//     segment are the left child.  Finished leaves carry key (index, 0) and keep their place;
//   * node monopoles bottom-up (leaf sums sequentially in tree order, parents = sum of children: deterministic),
//     Bmax by ordered-integer atomicMax along every particle's root-to-leaf path, cR2max = 4/3 Bmax^2 / theta^2;
//   * walk: one warp per 32 particles consecutive in tree order; the warp walks the union of its lanes' paths without a
//     stack (64-byte node records with a "next" link), each lane applying MarkCell's opening test r2 < cR2max in the
//     reference's operation order and accumulating monopole and leaf terms in fp64 in the reference's visiting order.
// Node MEMBERSHIP equals the reference's whenever the coordinates along a split dimension are distinct at the median
// (the reference's quickselect leaves the choice among exactly equal coordinates to its swap sequence; here the tie
// goes to the lower member index) -- see DESIGN.md "tree parity".
#include "groups.cuh"
#include "ptree.cuh"
#include "qsel.cuh"
#include "rsqrt.cuh"
#include "ufhash.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>

namespace vrgpu {

namespace {

constexpr uint32_t NOCHILD = 0xffffffffu;

__device__ __forceinline__ unsigned long long pt_ord(double d)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(d);
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double pt_unord(unsigned long long u)
{
    u = (u & 0x8000000000000000ull) ? (u & 0x7fffffffffffffffull) : ~u;
    return __longlong_as_double((long long)u);
}

// ---- selection of the large groups ------------------------------------------------------------------
__global__ void k_pt_flag(const uint32_t *__restrict__ gid, const uint32_t *__restrict__ off, uint32_t ntot, long long pp,
                          uint32_t *__restrict__ flags)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot) return;
    const uint32_t g = gid[s];
    flags[s] = (long long)(off[g + 1] - off[g]) > pp ? 1u : 0u;
}
__global__ void k_pt_compact(const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl, uint32_t ntot,
                             uint32_t *__restrict__ slot_of)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < ntot && flags[s]) slot_of[excl[s]] = s;
}
__global__ void k_pt_gflag(const uint32_t *__restrict__ off, uint32_t G, long long pp, uint32_t *__restrict__ gflag)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g > G) return;
    gflag[g] = (g >= 1 && (long long)(off[g + 1] - off[g]) > pp) ? 1u : 0u;
}
// root node r (= rank of the group among the large ones) covers the group's compacted range
__global__ void k_pt_roots(const uint32_t *__restrict__ off, const uint32_t *__restrict__ gflag, const uint32_t *__restrict__ gexcl,
                           const uint32_t *__restrict__ sexcl, uint32_t G, uint32_t ntot, uint32_t *__restrict__ nstart,
                           uint32_t *__restrict__ nend, uint32_t *__restrict__ nleft, uint32_t *__restrict__ frontier)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g > G || !gflag[g]) return;
    const uint32_t r = gexcl[g];
    const uint32_t s0 = off[g];
    const uint32_t c0 = sexcl[s0];
    nstart[r] = c0;
    nend[r] = c0 + (off[g + 1] - s0);
    nleft[r] = NOCHILD;
    frontier[r] = r;
}

// ---- ranks along each axis --------------------------------------------------------------------------
__global__ void k_pt_coord_keys(const double *__restrict__ c, const uint32_t *__restrict__ slot_of, uint32_t nL,
                                uint64_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nL) return;
    keys[p] = pt_ord(c[slot_of[p]]);
    vals[p] = p;
}
__global__ void k_pt_rank(const uint32_t *__restrict__ vals, uint32_t nL, uint32_t *__restrict__ rank)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < nL) rank[vals[p]] = p;
}
__global__ void k_pt_iota(uint32_t *__restrict__ a, uint32_t n)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) a[p] = p;
}

// ---- one level ----------------------------------------------------------------------------------------
__global__ void k_pt_heads(const uint32_t *__restrict__ frontier, uint32_t nF, const uint32_t *__restrict__ nstart, uint32_t nAct,
                           uint32_t *__restrict__ heads)
{
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nF) return;
    const uint32_t s = nstart[frontier[f]];
    if (s < nAct) heads[s] = 1u;
}
// fidx[pos] = index of the frontier entry that contains pos (exclusive scan of heads, + head - 1)
__global__ void k_pt_fidx(const uint32_t *__restrict__ heads, const uint32_t *__restrict__ excl, uint32_t nL, uint32_t *__restrict__ fidx)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < nL) fidx[p] = excl[p] + heads[p] - 1u;
}

struct PtCols { const double *x, *y, *z; };

// min / max per axis of every frontier node that will split (size > bucket)
__global__ void __launch_bounds__(256) k_pt_minmax(PtCols C, const uint32_t *__restrict__ slot_of, const uint32_t *__restrict__ perm,
                                                   const uint32_t *__restrict__ fidx, const uint32_t *__restrict__ frontier,
                                                   const uint32_t *__restrict__ nstart, const uint32_t *__restrict__ nend, uint32_t nL,
                                                   uint32_t bucket, unsigned long long *__restrict__ mm /* 6 per frontier entry */)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t f = NOCHILD;
    double v[3] = {0, 0, 0};
    if (p < nL) {
        f = fidx[p];
        const uint32_t nd = frontier[f];
        if (nend[nd] - nstart[nd] > bucket) {
            const uint32_t s = slot_of[perm[p]];
            v[0] = C.x[s]; v[1] = C.y[s]; v[2] = C.z[s];
        } else f = NOCHILD;
    }
    const uint32_t f0 = __shfl_sync(0xffffffffu, f, 0);
    if (__all_sync(0xffffffffu, f == f0)) {
        if (f0 == NOCHILD) return;
        for (int k = 0; k < 3; k++) {
            double mn = v[k], mx = v[k];
            for (int o = 16; o; o >>= 1) {
                mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
                mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            }
            if ((threadIdx.x & 31) == 0) { atomicMin(&mm[6 * (size_t)f0 + k], pt_ord(mn)); atomicMax(&mm[6 * (size_t)f0 + 3 + k], pt_ord(mx)); }
        }
    } else if (f != NOCHILD) {
        for (int k = 0; k < 3; k++) { atomicMin(&mm[6 * (size_t)f + k], pt_ord(v[k])); atomicMax(&mm[6 * (size_t)f + 3 + k], pt_ord(v[k])); }
    }
}

// split flag + split dimension (KDTree.cxx:438-507: largest spread, the first dimension wins a tie)
__global__ void k_pt_choose(const uint32_t *__restrict__ frontier, uint32_t nF, const uint32_t *__restrict__ nstart,
                            const uint32_t *__restrict__ nend, uint32_t bucket, const unsigned long long *__restrict__ mm,
                            uint32_t *__restrict__ splitflag, uint32_t *__restrict__ splitdim)
{
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nF) return;
    const uint32_t nd = frontier[f];
    const bool sp = nend[nd] - nstart[nd] > bucket;
    splitflag[f] = sp ? 1u : 0u;
    uint32_t d = 0;
    if (sp) {
        double best = __dsub_rn(pt_unord(mm[6 * (size_t)f + 3]), pt_unord(mm[6 * (size_t)f]));
        for (int k = 1; k < 3; k++) {
            double s = __dsub_rn(pt_unord(mm[6 * (size_t)f + 3 + k]), pt_unord(mm[6 * (size_t)f + k]));
            if (s > best) { best = s; d = k; }
        }
    }
    splitdim[f] = d;
}

__global__ void k_pt_level_keys(const uint32_t *__restrict__ perm, const uint32_t *__restrict__ fidx, const uint32_t *__restrict__ splitflag,
                                const uint32_t *__restrict__ splitdim, const uint32_t *__restrict__ rank /* 3 x nL */, uint32_t nL,
                                uint32_t nAct, int rb, uint64_t *__restrict__ keys)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nAct) return;
    const uint32_t f = fidx[p];
    uint32_t lo = 0;
    if (splitflag[f]) lo = rank[(size_t)splitdim[f] * nL + perm[p]];
    keys[p] = ((uint64_t)f << rb) | lo;
}

// children of the splitting nodes + next frontier (ordered by start)
__global__ void k_pt_children(const uint32_t *__restrict__ frontier, uint32_t nF, const uint32_t *__restrict__ splitflag,
                              const uint32_t *__restrict__ sexcl, uint32_t node_base, uint32_t *__restrict__ nstart,
                              uint32_t *__restrict__ nend, uint32_t *__restrict__ nleft, uint32_t *__restrict__ next_frontier)
{
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nF) return;
    const uint32_t nd = frontier[f];
    const uint32_t k = sexcl[f];
    if (!splitflag[f]) { next_frontier[f + k] = nd; return; }
    const uint32_t s = nstart[nd], e = nend[nd];
    const uint32_t mid = s + (e - s - 1) / 2 + 1; // left = [s, k+1) with k = s + (size-1)/2
    const uint32_t l = node_base + 2 * k, r = l + 1;
    nleft[nd] = l;                                // right child = left + 1
    nstart[l] = s; nend[l] = mid; nleft[l] = NOCHILD;
    nstart[r] = mid; nend[r] = e; nleft[r] = NOCHILD;
    next_frontier[f + k] = l;
    next_frontier[f + k + 1] = r;
}

// ---- VRGPU_TREE_TIES=1: the reference's quickselect arrangement instead of the sort (qsel.cuh) ---------------------
__global__ void k_qs_init(qsel::Level L)
{
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < L.nF) qsel::init_thread(L, f);
}
__global__ void k_qs_begin(qsel::Level L)
{
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < L.nF) qsel::begin_thread(L, f);
}
__global__ void k_qs_flag(qsel::Level L)
{
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w < L.W + 1u) qsel::flag_thread(L, w);
}
__global__ void k_qs_scatter(qsel::Level L)
{
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w < L.W) qsel::scatter_thread(L, w);
}
__global__ void k_qs_match(qsel::Level L)
{
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w < L.W) qsel::match_thread(L, w);
}
__global__ void k_qs_end(qsel::Level L)
{
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < L.nF) qsel::end_thread(L, f);
}
// nodes of at most QS_SMALL particles: a copy of the node in shared memory, MedianPos itself on one thread
constexpr int QS_SMALL_THREADS = 64;
constexpr size_t QS_SMALL_SMEM = (size_t)qsel::QS_SMALL * (sizeof(double) + sizeof(uint32_t));
__global__ void __launch_bounds__(QS_SMALL_THREADS) k_qs_small(qsel::Level L)
{
    extern __shared__ __align__(8) unsigned char qs_smem[];
    double *val = reinterpret_cast<double *>(qs_smem);
    uint32_t *idx = reinterpret_cast<uint32_t *>(val + qsel::QS_SMALL);
    for (uint32_t f = blockIdx.x; f < L.nF; f += gridDim.x) {
        uint32_t s, e;
        if (!qsel::small_node(L, f, s, e)) continue; // the same answer for the whole CTA
        const uint32_t n = e - s, d = L.splitdim[f];
        for (uint32_t i = threadIdx.x; i < n; i += QS_SMALL_THREADS) {
            const uint32_t m = L.perm[s + i];
            idx[i] = m;
            val[i] = qsel::coord(L, d, m);
        }
        __syncthreads();
        if (threadIdx.x == 0) qsel::select_serial(val, idx, qsel::median_index(s, e) - s, n);
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < n; i += QS_SMALL_THREADS) L.perm[s + i] = idx[i];
        __syncthreads();
    }
}

} // namespace

namespace {

// ---- monopoles ------------------------------------------------------------------------------------------
struct PtNodes {
    uint32_t *start, *end, *left;
    double *sx, *sy, *sz, *sm;      // sums, then centre of mass in place
    unsigned long long *bmax;       // ordered-int image of Bmax
    double *r2max;
};

// leaves: sequential sums in tree order (unbind.cxx:1672-1675: cm += x*m, product rounded before the add)
__global__ void k_pt_leaf_sums(PtNodes N, uint32_t nnodes, const double *__restrict__ tx, const double *__restrict__ ty,
                               const double *__restrict__ tz, const double *__restrict__ tm)
{
    uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= nnodes || N.left[nd] != NOCHILD) return;
    double sx = 0, sy = 0, sz = 0, sm = 0;
    for (uint32_t p = N.start[nd]; p < N.end[nd]; p++) {
        const double m = tm[p];
        sx = __dadd_rn(sx, __dmul_rn(tx[p], m)); sy = __dadd_rn(sy, __dmul_rn(ty[p], m)); sz = __dadd_rn(sz, __dmul_rn(tz[p], m));
        sm = __dadd_rn(sm, m);
    }
    N.sx[nd] = sx; N.sy[nd] = sy; N.sz[nd] = sz; N.sm[nd] = sm;
}
// internal nodes of one level: sum of the two children
__global__ void k_pt_inner_sums(PtNodes N, uint32_t n0, uint32_t n1)
{
    uint32_t nd = n0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= n1) return;
    const uint32_t l = N.left[nd];
    if (l == NOCHILD) return;
    N.sx[nd] = __dadd_rn(N.sx[l], N.sx[l + 1]); N.sy[nd] = __dadd_rn(N.sy[l], N.sy[l + 1]);
    N.sz[nd] = __dadd_rn(N.sz[l], N.sz[l + 1]); N.sm[nd] = __dadd_rn(N.sm[l], N.sm[l + 1]);
}
__global__ void k_pt_cm(PtNodes N, uint32_t nnodes)
{
    uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= nnodes) return;
    const double m = N.sm[nd];
    N.sx[nd] = __ddiv_rn(N.sx[nd], m); N.sy[nd] = __ddiv_rn(N.sy[nd], m); N.sz[nd] = __ddiv_rn(N.sz[nd], m);
    N.bmax[nd] = 0ull;
}
// Bmax (unbind.cxx:1677-1683): max over the node's particles of |cm - x|; every particle visits its root-to-leaf path
__global__ void __launch_bounds__(256) k_pt_bmax(PtNodes N, const uint32_t *__restrict__ rootidx, uint32_t nL, const double *__restrict__ tx,
                                                 const double *__restrict__ ty, const double *__restrict__ tz)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = p < nL;
    uint32_t nd = valid ? rootidx[p] : NOCHILD;
    double x = 0, y = 0, z = 0;
    if (valid) { x = tx[p]; y = ty[p]; z = tz[p]; }
    while (__any_sync(0xffffffffu, nd != NOCHILD)) {
        double l = -1.0;
        if (nd != NOCHILD) {
            const double dx = __dsub_rn(N.sx[nd], x), dy = __dsub_rn(N.sy[nd], y), dz = __dsub_rn(N.sz[nd], z);
            l = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
        }
        const uint32_t n0 = __shfl_sync(0xffffffffu, nd, 0);
        if (__all_sync(0xffffffffu, nd == n0)) {
            for (int o = 16; o; o >>= 1) l = fmax(l, __shfl_xor_sync(0xffffffffu, l, o));
            if ((threadIdx.x & 31) == 0) atomicMax(&N.bmax[n0], pt_ord(l));
        } else if (nd != NOCHILD) atomicMax(&N.bmax[nd], pt_ord(l));
        if (nd != NOCHILD) {
            const uint32_t lc = N.left[nd];
            nd = lc == NOCHILD ? NOCHILD : (p < N.end[lc] ? lc : lc + 1);
        }
    }
}
__global__ void k_pt_r2max(PtNodes N, uint32_t nnodes, double theta)
{
    uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= nnodes) return;
    const double xdiff = pt_unord(N.bmax[nd]);
    // cR2max = 4.0/3.0*xdiff*xdiff/(theta*theta)   (unbind.cxx:1684, evaluated left to right)
    N.r2max[nd] = __ddiv_rn(__dmul_rn(__dmul_rn(4.0 / 3.0, xdiff), xdiff), __dmul_rn(theta, theta));
}


// ---- walk (MarkCell, unbind.cxx:24-44 + the sums at :1710-1723) ---------------------------------------------
// One 64-byte record per node, and a "next" link (the node that follows the whole subtree in the reference's
// left-before-right recursion) so the walk needs no stack.
struct __align__(16) PtNode64 {
    double cx, cy, cz, r2max;
    double sm;
    uint32_t left, start;
    uint32_t end, next;
    uint32_t pad[2];
};
static_assert(sizeof(PtNode64) == 64, "node record is 4 x 16 bytes");

__global__ void k_pt_next(const uint32_t *__restrict__ left, uint32_t *__restrict__ next, uint32_t n0, uint32_t n1)
{
    const uint32_t nd = n0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= n1) return;
    const uint32_t l = left[nd];
    if (l == NOCHILD) return;
    next[l] = l + 1;
    next[l + 1] = next[nd];
}
__global__ void k_pt_pack(PtNodes N, const uint32_t *__restrict__ next, uint32_t nnodes, PtNode64 *__restrict__ out)
{
    const uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= nnodes) return;
    PtNode64 r;
    r.cx = N.sx[nd]; r.cy = N.sy[nd]; r.cz = N.sz[nd]; r.r2max = N.r2max[nd]; r.sm = N.sm[nd];
    r.left = N.left[nd]; r.start = N.start[nd]; r.end = N.end[nd]; r.next = next[nd];
    r.pad[0] = r.pad[1] = 0;
    out[nd] = r;
}

// One warp walks the tree ONCE for its 32 particles (consecutive in tree order, so their walks overlap almost entirely):
// the warp descends wherever any of its lanes opens a node; a lane that accepts a node's monopole idles until the walk
// leaves that subtree (cur == the node's next link).  Every lane therefore sees exactly the nodes, in exactly the order,
// of the reference's per-particle recursion, and node records are fetched once per warp instead of once per lane.
constexpr uint32_t PT_ACTIVE = 0xfffffffeu;
// Measured and dropped (profiles/r02a_probe.md): prefetch.global.L1 of both possible successors of a node -- 10 % slower.
__global__ void __launch_bounds__(128) k_pt_walk(const PtNode64 *__restrict__ nodes, const uint32_t *__restrict__ rootidx, uint32_t nL,
                                                 const double *__restrict__ tx, const double *__restrict__ ty, const double *__restrict__ tz,
                                                 const double *__restrict__ tm, double G, double eps2, const uint32_t *__restrict__ perm,
                                                 const uint32_t *__restrict__ slot_of, double *__restrict__ phi,
                                                 unsigned long long *__restrict__ ninter)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = p < nL;
    const uint32_t myroot = valid ? rootidx[p] : NOCHILD;
    double x = 0, y = 0, z = 0, mj = 0;
    if (valid) { x = tx[p]; y = ty[p]; z = tz[p]; mj = tm[p]; }
    double acc = 0.0;
    unsigned long long cnt = 0;
    unsigned todo = __ballot_sync(0xffffffffu, valid);
    while (todo) {
        const uint32_t root = __shfl_sync(0xffffffffu, myroot, __ffs(todo) - 1);
        const bool mine = valid && myroot == root;
        todo &= ~__ballot_sync(0xffffffffu, mine);
        uint32_t cur = root, wait = PT_ACTIVE;
        while (cur != NOCHILD) {
            const double2 *q = reinterpret_cast<const double2 *>(nodes + cur);
            const double2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            const uint2 d = __ldg(reinterpret_cast<const uint2 *>(q + 3));
            const uint32_t left = (uint32_t)(__double_as_longlong(c.y) & 0xffffffffll);
            const uint32_t start = (uint32_t)((unsigned long long)__double_as_longlong(c.y) >> 32);
            const uint32_t end = d.x, next = d.y;
            if (wait == cur) wait = PT_ACTIVE;
            const bool act = mine && wait == PT_ACTIVE;
            const double dx = __dsub_rn(a.x, x), dy = __dsub_rn(a.y, y), dz = __dsub_rn(b.x, z);
            const double r2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
            const bool open = act && r2 < b.y;
            if (act && !open) {
                acc -= mj * (c.x * rsqrt_refined(r2 + eps2));
                cnt++;
                wait = next;
            }
            if (!__any_sync(0xffffffffu, open)) { cur = next; continue; }
            if (left != NOCHILD) { cur = left; continue; }
            for (uint32_t l = start; l < end; l++) {
                const double lx = __ldg(tx + l), ly = __ldg(ty + l), lz = __ldg(tz + l), lm = __ldg(tm + l);
                if (open && l != p) {
                    const double ax = __dsub_rn(x, lx), ay = __dsub_rn(y, ly), az = __dsub_rn(z, lz);
                    const double q2 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(ax, ax), __dmul_rn(ay, ay)), __dmul_rn(az, az)), eps2);
                    acc -= (mj * lm) * rsqrt_refined(q2);
                    cnt++;
                }
            }
            cur = next;
        }
    }
    if (valid) phi[slot_of[perm[p]]] = acc * G;
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(ninter, cnt);
}

__global__ void k_pt_gather_tree(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                                 const double *__restrict__ m, const uint32_t *__restrict__ perm, const uint32_t *__restrict__ slot_of, uint32_t nL,
                                 double *__restrict__ tx, double *__restrict__ ty, double *__restrict__ tz, double *__restrict__ tm)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nL) return;
    const uint32_t s = slot_of[perm[p]];
    tx[p] = x[s]; ty[p] = y[s]; tz[p] = z[s]; tm[p] = m[s];
}

} // namespace

namespace {
__global__ void k_pt_mm_init(unsigned long long *__restrict__ mm, uint32_t nF)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 6 * nF) mm[i] = (i % 6) < 3 ? ~0ull : 0ull;
}
} // namespace

// The build shared by the potentials (below) and the local velocity density (veldens.cu): members of the groups above
// pp_threshold in tree order (t.perm over compact member indices, t.slot_of to the group-major columns) and the nodes over
// them, numbered level by level (right child = left + 1).  False when no group qualifies.
bool pt_build(vrgpu_ctx *ctx, const GroupSet &gs, int bucket, long long pp_threshold, bool exact_ties, PtTree &t)
{
    t.nL = t.nLG = t.nnodes = 0;
    t.level_off.clear();
    if (gs.ntot == 0 || gs.ngroups == 0 || pp_threshold < 0) return false;
    VR_REQUIRE(bucket >= 1, VRGPU_EINVAL, "BucketSize must be >= 1");
    const int T = 256;
    const uint32_t ntot = (uint32_t)gs.ntot, NG = (uint32_t)gs.ngroups;
    DBuf<uint32_t> &slot_of = t.slot_of, &perm = t.perm, &nstart = t.nstart, &nend = t.nend, &nleft = t.nleft, &rootidx = t.rootidx;
    std::vector<uint32_t> &level_off = t.level_off;

    // ---- large groups -> compact positions ----
    DBuf<uint32_t> flags(ctx, ntot), sexcl(ctx, ntot), total(ctx, 2), gflag(ctx, (size_t)NG + 1), gexcl(ctx, (size_t)NG + 1);
    VR_LAUNCH(ctx, k_pt_flag, grid_for(ntot, T), T, 0, gs.gid.p, gs.off.p, ntot, pp_threshold, flags.p);
    exclusive_scan_u32(ctx, flags.p, sexcl.p, ntot, total.p);
    VR_LAUNCH(ctx, k_pt_gflag, grid_for((size_t)NG + 1, T), T, 0, gs.off.p, NG, pp_threshold, gflag.p);
    exclusive_scan_u32(ctx, gflag.p, gexcl.p, (size_t)NG + 1, total.p + 1);
    uint32_t h2[2];
    VR_CUDA(cudaMemcpyAsync(h2, total.p, sizeof(h2), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint32_t nL = h2[0], nLG = h2[1];
    if (nL == 0) return false;
    slot_of.alloc(ctx, nL);
    VR_LAUNCH(ctx, k_pt_compact, grid_for(ntot, T), T, 0, flags.p, sexcl.p, ntot, slot_of.p);

    // ---- ranks along the three axes (ties: lower member first); exact_ties arranges every node with the reference's own
    // selection instead (qsel.cuh) and needs none ----
    perm.alloc(ctx, nL);
    DBuf<uint32_t> rank(ctx, exact_ties ? 1 : 3 * (size_t)nL), vtmp(ctx, exact_ties ? 1 : nL);
    DBuf<uint64_t> keys(ctx, exact_ties ? 1 : nL), ktmp(ctx, exact_ties ? 1 : nL);
    if (!exact_ties) {
        KScope k(ctx, "ptree.ranks");
        const double *cols[3] = {gs.x.p, gs.y.p, gs.z.p};
        for (int d = 0; d < 3; d++) {
            VR_LAUNCH(ctx, k_pt_coord_keys, grid_for(nL, T), T, 0, cols[d], slot_of.p, nL, keys.p, perm.p);
            radix_sort_pairs(ctx, keys.p, perm.p, ktmp.p, vtmp.p, nL, 64);
            VR_LAUNCH(ctx, k_pt_rank, grid_for(nL, T), T, 0, perm.p, nL, rank.p + (size_t)d * nL);
        }
    }

    // ---- nodes ----
    const uint32_t minleaf = std::max(1u, ((uint32_t)bucket + 1u) / 2u);
    const size_t cap = 2 * ((size_t)nL / minleaf + 1) + nLG + 16;
    nstart.alloc(ctx, cap); nend.alloc(ctx, cap); nleft.alloc(ctx, cap);
    rootidx.alloc(ctx, nL);
    DBuf<uint32_t> frontier(ctx, cap), frontier2(ctx, cap), splitflag(ctx, cap), splitdim(ctx, cap), fexcl(ctx, cap);
    DBuf<uint32_t> heads(ctx, nL), hexcl(ctx, nL), fidx(ctx, nL);
    DBuf<unsigned long long> mm(ctx, 6 * cap);
    // selection state of the exact-ties variant
    const size_t qcap = exact_ties ? cap : 1, qn = exact_ties ? (size_t)nL + 1 : 1;
    DBuf<uint32_t> qlo(ctx, qcap), qhi(ctx, qcap), qnp(ctx, qcap), qlen(ctx, qcap), qwoff(ctx, qcap), qtotal(ctx, 1);
    DBuf<double> qpiv(ctx, qcap);
    DBuf<unsigned long long> qflags(ctx, qn), qscan(ctx, qn);
    DBuf<uint32_t> qwf(ctx, qn), qA(ctx, qn), qB(ctx, qn);
    VR_LAUNCH(ctx, k_pt_roots, grid_for((size_t)NG + 1, T), T, 0, gs.off.p, gflag.p, gexcl.p, sexcl.p, NG, ntot, nstart.p, nend.p, nleft.p, frontier.p);
    VR_LAUNCH(ctx, k_pt_iota, grid_for(nL, T), T, 0, perm.p, nL);
    uint32_t nF = nLG, nnodes = nLG;
    level_off.assign({0u, nLG});
    // A group's largest node at level l holds ceil(n / 2^l) particles (median splits), so the group is finished once that
    // is <= bucket: the per-particle passes of a level only cover positions below the end of the last unfinished group
    // (groups come in size order from the search stages, so this prefix shrinks level by level).
    std::vector<uint32_t> gstart(nLG), gend(nLG);
    VR_CUDA(cudaMemcpyAsync(gstart.data(), nstart.p, sizeof(uint32_t) * nLG, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(gend.data(), nend.p, sizeof(uint32_t) * nLG, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    auto active_prefix = [&](int level) {
        uint32_t top = 0;
        for (uint32_t r = 0; r < nLG; r++) {
            const uint64_t n = gend[r] - gstart[r];
            const uint64_t big = level < 40 ? ((n + (1ull << level) - 1) >> level) : 1;
            if (big > (uint64_t)bucket) top = std::max(top, gend[r]);
        }
        return top;
    };
    const int rb = bits_for_u64v(nL > 1 ? nL - 1 : 1);
    PtCols C{gs.x.p, gs.y.p, gs.z.p};
    uint32_t *fr = frontier.p, *fr2 = frontier2.p;
    {
        KScope kb(ctx, "ptree.build");
        for (int level = 0; level < 64; level++) {
            const uint32_t nAct = level == 0 ? nL : active_prefix(level);
            if (nAct == 0) break;
            VR_CUDA(cudaMemsetAsync(heads.p, 0, sizeof(uint32_t) * nAct, ctx->stream));
            VR_LAUNCH(ctx, k_pt_heads, grid_for(nF, T), T, 0, fr, nF, nstart.p, nAct, heads.p);
            exclusive_scan_u32(ctx, heads.p, hexcl.p, nAct, nullptr);
            VR_LAUNCH(ctx, k_pt_fidx, grid_for(nAct, T), T, 0, heads.p, hexcl.p, nAct, fidx.p);
            if (level == 0) VR_CUDA(cudaMemcpyAsync(rootidx.p, fidx.p, sizeof(uint32_t) * nL, cudaMemcpyDeviceToDevice, ctx->stream));
            VR_LAUNCH(ctx, k_pt_mm_init, grid_for((size_t)6 * nF, T), T, 0, mm.p, nF);
            VR_LAUNCH(ctx, k_pt_minmax, grid_for(nAct, T), T, 0, C, slot_of.p, perm.p, fidx.p, fr, nstart.p, nend.p, nAct, (uint32_t)bucket, mm.p);
            VR_LAUNCH(ctx, k_pt_choose, grid_for(nF, T), T, 0, fr, nF, nstart.p, nend.p, (uint32_t)bucket, mm.p, splitflag.p, splitdim.p);
            exclusive_scan_u32(ctx, splitflag.p, fexcl.p, nF, total.p);
            uint32_t nsplit = 0;
            VR_CUDA(cudaMemcpyAsync(&nsplit, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
            if (nsplit == 0) break;
            VR_REQUIRE((size_t)nnodes + 2 * (size_t)nsplit <= cap, VRGPU_ECUDA, "kd-tree node capacity exceeded");
            if (exact_ties) {
                qsel::Level L{gs.x.p, gs.y.p, gs.z.p, slot_of.p, perm.p, fr, nstart.p, nend.p, splitflag.p, splitdim.p, qlo.p, qhi.p,
                              qnp.p, qlen.p, qwoff.p, qpiv.p, qwf.p, qflags.p, qscan.p, qA.p, qB.p, nF, 0u};
                VR_LAUNCH(ctx, k_qs_init, grid_for(nF, T), T, 0, L);
                VR_LAUNCH(ctx, k_qs_small, (unsigned)std::min<size_t>(nF, (size_t)148 * 64), QS_SMALL_THREADS, QS_SMALL_SMEM, L);
                for (;;) { // one iteration of MedianPos for every node above QS_SMALL that is not finished
                    exclusive_scan_u32(ctx, qlen.p, qwoff.p, nF, qtotal.p);
                    VR_CUDA(cudaMemcpyAsync(&L.W, qtotal.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
                    VR_CUDA(cudaStreamSynchronize(ctx->stream));
                    if (L.W == 0) break;
                    VR_LAUNCH(ctx, k_qs_begin, grid_for(nF, T), T, 0, L);
                    VR_LAUNCH(ctx, k_qs_flag, grid_for((size_t)L.W + 1, T), T, 0, L);
                    exclusive_scan_u64(ctx, qflags.p, qscan.p, (size_t)L.W + 1, nullptr);
                    VR_LAUNCH(ctx, k_qs_scatter, grid_for(L.W, T), T, 0, L);
                    VR_LAUNCH(ctx, k_qs_match, grid_for(L.W, T), T, 0, L);
                    VR_LAUNCH(ctx, k_qs_end, grid_for(nF, T), T, 0, L);
                }
            } else {
                VR_LAUNCH(ctx, k_pt_level_keys, grid_for(nAct, T), T, 0, perm.p, fidx.p, splitflag.p, splitdim.p, rank.p, nL, nAct, rb, keys.p);
                radix_sort_pairs(ctx, keys.p, perm.p, ktmp.p, vtmp.p, nAct, rb + bits_for_u64v(nF > 1 ? nF - 1 : 1));
            }
            VR_LAUNCH(ctx, k_pt_children, grid_for(nF, T), T, 0, fr, nF, splitflag.p, fexcl.p, nnodes, nstart.p, nend.p, nleft.p, fr2);
            nnodes += 2 * nsplit;
            nF += nsplit;
            level_off.push_back(nnodes);
            std::swap(fr, fr2);
        }
    }
    t.nL = nL;
    t.nLG = nLG;
    t.nnodes = nnodes;
    return true;
}

void potential_tree(vrgpu_ctx *ctx, const GroupSet &gs, double G, double eps, double theta, int bucket, long long pp_threshold, double *phi)
{
    if (gs.ntot == 0 || gs.ngroups == 0 || pp_threshold < 0) return;
    VR_REQUIRE(theta > 0, VRGPU_EINVAL, "TreeThetaOpen must be positive");
    // VRGPU_TREE_TIES=1 (read per call): arrange every node with the reference's own selection, see qsel.cuh
    const bool exact_ties = [] { const char *e = getenv("VRGPU_TREE_TIES"); return e && atoi(e) == 1; }();
    PtTree t;
    if (!pt_build(ctx, gs, bucket, pp_threshold, exact_ties, t)) return;
    const int T = 256;
    const uint32_t nL = t.nL, nLG = t.nLG, nnodes = t.nnodes;
    DBuf<uint32_t> &slot_of = t.slot_of, &perm = t.perm, &nstart = t.nstart, &nend = t.nend, &nleft = t.nleft, &rootidx = t.rootidx;
    const std::vector<uint32_t> &level_off = t.level_off;

    // ---- tree-ordered columns, monopoles, opening radii ----
    DBuf<double> tx(ctx, nL), ty(ctx, nL), tz(ctx, nL), tm(ctx, nL);
    DBuf<double> sx(ctx, nnodes), sy(ctx, nnodes), sz(ctx, nnodes), sm(ctx, nnodes), r2max(ctx, nnodes);
    DBuf<unsigned long long> bmax(ctx, nnodes);
    PtNodes N{nstart.p, nend.p, nleft.p, sx.p, sy.p, sz.p, sm.p, bmax.p, r2max.p};
    {
        KScope k(ctx, "ptree.moments");
        VR_LAUNCH(ctx, k_pt_gather_tree, grid_for(nL, T), T, 0, gs.x.p, gs.y.p, gs.z.p, gs.m.p, perm.p, slot_of.p, nL, tx.p, ty.p, tz.p, tm.p);
        VR_LAUNCH(ctx, k_pt_leaf_sums, grid_for(nnodes, T), T, 0, N, nnodes, tx.p, ty.p, tz.p, tm.p);
        for (size_t l = level_off.size() - 1; l-- > 0;) {
            const uint32_t n0 = level_off[l], n1 = level_off[l + 1];
            if (n1 > n0) VR_LAUNCH(ctx, k_pt_inner_sums, grid_for(n1 - n0, T), T, 0, N, n0, n1);
        }
        VR_LAUNCH(ctx, k_pt_cm, grid_for(nnodes, T), T, 0, N, nnodes);
        VR_LAUNCH(ctx, k_pt_bmax, grid_for(nL, T), T, 0, N, rootidx.p, nL, tx.p, ty.p, tz.p);
        VR_LAUNCH(ctx, k_pt_r2max, grid_for(nnodes, T), T, 0, N, nnodes, theta);
    }
    // ---- walk ----
    DBuf<unsigned long long> ninter(ctx, 1);
    DBuf<PtNode64> packed(ctx, nnodes);
    VR_CUDA(cudaMemsetAsync(ninter.p, 0, sizeof(unsigned long long), ctx->stream));
    {
        KScope k(ctx, "ptree.walk");
        DBuf<uint32_t> next(ctx, nnodes);
        VR_CUDA(cudaMemsetAsync(next.p, 0xff, sizeof(uint32_t) * nLG, ctx->stream)); // roots: nothing follows
        for (size_t l = 0; l + 1 < level_off.size(); l++) {
            const uint32_t n0 = level_off[l], n1 = level_off[l + 1];
            if (n1 > n0) VR_LAUNCH(ctx, k_pt_next, grid_for(n1 - n0, T), T, 0, nleft.p, next.p, n0, n1);
        }
        VR_LAUNCH(ctx, k_pt_pack, grid_for(nnodes, T), T, 0, N, next.p, nnodes, packed.p);
        VR_LAUNCH(ctx, k_pt_walk, grid_for(nL, 128), 128, 0, packed.p, rootidx.p, nL, tx.p, ty.p, tz.p, tm.p, G, eps * eps, perm.p, slot_of.p,
                  phi, ninter.p);
    }
    unsigned long long hi = 0;
    VR_CUDA(cudaMemcpyAsync(&hi, ninter.p, sizeof(hi), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->stats.n_tree_interactions += (long long)hi;
    ctx->stats.n_tree_nodes += (long long)nnodes;
}

} // namespace vrgpu
