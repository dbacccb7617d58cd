// velociraptor-stf_b200/csrc/ptree.cu -- kd-tree monopole ("Barnes-Hut") potentials for large groups.
//
// Replaces Potential() (src/unbind.cxx:1607-1741) with its helpers GetNodeList (:13-21) and MarkCell (:24-44) and the
// kd-tree it builds (NBody::KDTree::BuildNodes, NBodylib/src/KDTree/KDTree.cxx:438-507: split dimension = largest
// spread (first wins), median index k = start + (size-1)/2, left child = the k-start+1 smallest along that dimension,
// leaf when size <= BucketSize).
//
// Device design: every group above the direct-sum threshold is built in the SAME launches.
//   * nodes: the tree's shape follows from the group sizes alone (left child = ceil(size / 2) particles), so the node
//     arrays are laid down first, level by level, without touching a particle;
//   * arrangement: MedianPos itself (the reference's quickselect, swap for swap -- see "the reference's arrangement"
//     below): nodes above 2048 particles by one CTA each in global memory, every subtree of at most 2048 particles by one
//     CTA in shared memory.  Node membership AND particle order are the reference's for every input, tied coordinates
//     (float32 snapshots) included;
//   * node monopoles bottom-up (leaf sums sequentially in tree order, parents = sum of children: deterministic),
//     Bmax by ordered-integer atomicMax along every particle's root-to-leaf path, cR2max = 4/3 Bmax^2 / theta^2;
//   * walk: one warp per 32 particles consecutive in tree order; the warp walks the union of its lanes' paths without a
//     stack (64-byte node records with a "next" link), each lane applying MarkCell's opening test r2 < cR2max in the
//     reference's operation order and accumulating monopole and leaf terms in fp64 in the reference's visiting order.
#include "groups.cuh"
#include "ptree.cuh"
#include "rsqrt.cuh"
#include "seqblock.cuh"
#include "ufhash.cuh"

#include <cooperative_groups.h>

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

__global__ void k_pt_iota(uint32_t *__restrict__ a, uint32_t n)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) a[p] = p;
}

// ---- tree position -> root node (head flags + scan) ---------------------------------------------------------
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

// ---- the reference's arrangement: MedianPos itself (KDTree.cxx:304-337) on every node --------------------------------
// BuildNodes splits a node with a quickselect (Hoare partition about the element AT the median index).  Which of several
// particles with exactly equal split coordinate go left, and in which order every particle is handed to the next level,
// follows from its swap sequence; float32 snapshots have such ties at most levels of a big group, and node membership
// decides which cells the walk opens.  So the build runs MedianPos, not a sort.  One iteration on the positions [lo, hi)
// of a node (pivot index k, pivot value x) is data parallel (tests/test_hoare_model.py pins this form against the literal
// loop):
//     swap p[k] <-> p[hi-1]
//     A = positions of [lo, hi-1) with value >= x, ascending (where the upward scan stops)
//     B = positions of [lo, hi-1) with value <= x, descending (where the downward scan stops)
//     M = number of m with A[m] < B[m] (a prefix: A increases, B decreases);  swap p[A[m]] <-> p[B[m]] for m < M
//     i* = A[M] if it exists and (M == 0 or A[M] < B[M-1]), else B[M-1] if M > 0, else hi-1;  swap p[i*] <-> p[hi-1]
//     hi = i* if i* >= k;  lo = i* + 1 if i* <= k;  repeat while lo + 1 < hi
// The tree SHAPE is a function of the group sizes alone (left child = ceil(size / 2) particles), so the node arrays are laid
// down first; then every node above KD_SMALL particles is arranged by one CTA in global memory, level by level, and every
// subtree of at most KD_SMALL particles is finished by one CTA in shared memory (one warp per node, sub-level by sub-level).
// Particles travel as 32-byte records (x, y, z, member), so a scan reads one sector per particle and a swap moves two.
constexpr uint32_t KD_SMALL = 2048;
struct __align__(16) KdRec { double x, y, z; uint32_t member, pad; };
static_assert(sizeof(KdRec) == 32, "one 32-byte sector per particle");

__global__ void k_kd_init_rec(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                              const uint32_t *__restrict__ slot_of, uint32_t nL, KdRec *__restrict__ rec)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nL) return;
    const uint32_t s = slot_of[p];
    KdRec r;
    r.x = x[s]; r.y = y[s]; r.z = z[s]; r.member = p; r.pad = 0;
    rec[p] = r;
}
__global__ void k_kd_perm(const KdRec *__restrict__ rec, uint32_t nL, uint32_t *__restrict__ perm)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < nL) perm[p] = rec[p].member;
}

// skeleton, one level: which nodes split; the nodes above KD_SMALL go to this level's list of CTA-arranged nodes, roots that
// fit shared memory to the list of subtrees
__global__ void k_kd_level_flags(const uint32_t *__restrict__ nstart, const uint32_t *__restrict__ nend, uint32_t n0, uint32_t n1, uint32_t bucket,
                                 int root_level, uint32_t *__restrict__ flags, uint32_t *__restrict__ big_list, uint32_t *__restrict__ sub_list,
                                 uint32_t *__restrict__ counters /* [0] big of this level, [1] subtrees */)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (n0 + i >= n1) return;
    const uint32_t nd = n0 + i, size = nend[nd] - nstart[nd];
    const bool split = size > bucket;
    flags[i] = split ? 1u : 0u;
    if (!split) return;
    if (size > KD_SMALL) big_list[atomicAdd(&counters[0], 1u)] = nd;
    else if (root_level) sub_list[atomicAdd(&counters[1], 1u)] = nd;
}
__global__ void k_kd_level_children(uint32_t n0, uint32_t n1, const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl, uint32_t bucket,
                                    uint32_t *__restrict__ nstart, uint32_t *__restrict__ nend, uint32_t *__restrict__ nleft,
                                    uint32_t *__restrict__ sub_list, uint32_t *__restrict__ counters)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (n0 + i >= n1 || !flags[i]) return;
    const uint32_t nd = n0 + i;
    const uint32_t s = nstart[nd], e = nend[nd];
    const uint32_t mid = s + (e - s - 1) / 2 + 1; // left = [s, k+1) with k = s + (size-1)/2
    const uint32_t l = n1 + 2 * excl[i], r = l + 1; // right child = left + 1
    nleft[nd] = l;
    nstart[l] = s; nend[l] = mid; nleft[l] = NOCHILD;
    nstart[r] = mid; nend[r] = e; nleft[r] = NOCHILD;
    if (e - s > KD_SMALL) { // children that fit shared memory start a subtree
        if (mid - s <= KD_SMALL && mid - s > bucket) sub_list[atomicAdd(&counters[1], 1u)] = l;
        if (e - mid <= KD_SMALL && e - mid > bucket) sub_list[atomicAdd(&counters[1], 1u)] = r;
    }
}

__device__ __forceinline__ void kd_swap_rec(KdRec *rec, uint32_t a, uint32_t b)
{
    uint4 *pa = reinterpret_cast<uint4 *>(rec + a), *pb = reinterpret_cast<uint4 *>(rec + b);
    const uint4 a0 = pa[0], a1 = pa[1], b0 = pb[0], b1 = pb[1];
    pa[0] = b0; pa[1] = b1; pb[0] = a0; pb[1] = a1;
}
__device__ __forceinline__ double kd_coord(const KdRec *rec, uint32_t p, int d) { return reinterpret_cast<const double *>(rec + p)[d]; }

constexpr int KB_NT = 1024, KB_ITEMS = 4, KB_TILE = KB_NT * KB_ITEMS;
// One CTA runs the reference's selection (MedianPos, KDTree.cxx:304-337) on the positions [lo, hi) of a node in global memory until
// position k holds the median: the body of k_kd_big, also the tail of k_kd_huge once the range is too short for the whole GPU.
__device__ void kd_cta_select(KdRec *rec, const uint32_t k, const int d, uint32_t lo, uint32_t hi, uint32_t *An, uint32_t *Bn)
{
    __shared__ uint32_t s_w[KB_NT / 32];
    __shared__ uint32_t s_tot, s_first, s_lo, s_hi;
    __shared__ double s_x;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    while (lo + 1 < hi) {
        if (tid == 0) {
            s_x = kd_coord(rec, k, d);
            if (k != hi - 1) kd_swap_rec(rec, k, hi - 1);
        }
        __syncthreads();
        const double x = s_x;
        const uint32_t top = hi - 1; // the scans cover [lo, top)
        // ---- the two lists (both ascending; B is read from its end) ----
        uint32_t nA = 0, nB = 0;
        double v[KB_ITEMS];
#pragma unroll
        for (int j = 0; j < KB_ITEMS; j++) {
            const uint32_t p = lo + tid * KB_ITEMS + j;
            v[j] = p < top ? kd_coord(rec, p, d) : 0.0;
        }
        for (uint32_t base = lo; base < top; base += KB_TILE) {
            double vn[KB_ITEMS];
            const uint32_t nbase = base + KB_TILE;
#pragma unroll
            for (int j = 0; j < KB_ITEMS; j++) { // the next tile's loads leave before this tile's scan
                const uint32_t p = nbase + tid * KB_ITEMS + j;
                vn[j] = (nbase < top && p < top) ? kd_coord(rec, p, d) : 0.0;
            }
            const uint32_t p0 = base + tid * KB_ITEMS;
            unsigned fa = 0, fb = 0;
#pragma unroll
            for (int j = 0; j < KB_ITEMS; j++)
                if (p0 + j < top) { fa |= (v[j] >= x ? 1u : 0u) << j; fb |= (v[j] <= x ? 1u : 0u) << j; }
            // exclusive scan over the CTA of (count A | count B << 16): at most 4096 each per tile
            const uint32_t mine = (uint32_t)__popc(fa) | ((uint32_t)__popc(fb) << 16);
            uint32_t incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= (unsigned)o) incl += y;
            }
            if (lane == 31) s_w[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                const uint32_t wv = s_w[lane];
                uint32_t wi = wv;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xffffffffu, wi, o);
                    if (lane >= (unsigned)o) wi += y;
                }
                s_w[lane] = wi - wv;
                if (lane == 31) s_tot = wi;
            }
            __syncthreads();
            const uint32_t excl = s_w[warp] + incl - mine, tot = s_tot;
            uint32_t ea = nA + (excl & 0xffffu), eb = nB + (excl >> 16);
#pragma unroll
            for (int j = 0; j < KB_ITEMS; j++) {
                if ((fa >> j) & 1u) An[ea++] = p0 + j;
                if ((fb >> j) & 1u) Bn[eb++] = p0 + j;
            }
            nA += tot & 0xffffu; nB += tot >> 16;
#pragma unroll
            for (int j = 0; j < KB_ITEMS; j++) v[j] = vn[j];
            __syncthreads(); // s_w / s_tot are rewritten by the next tile
        }
        // ---- M = first m with !(A[m] < B[nB-1-m]) (the predicate holds on a prefix): KB_NT-ary search ----
        const uint32_t mm = nA < nB ? nA : nB;
        uint32_t a = 0, b = mm;
        while (a < b) {
            const uint32_t step = (b - a + KB_NT - 1) / KB_NT;
            const uint32_t m = a + tid * step;
            const bool fail = m < b && !(An[m] < Bn[nB - 1 - m]);
            const unsigned bal = __ballot_sync(0xffffffffu, fail);
            if (lane == 0) s_w[warp] = bal ? warp * 32 + (uint32_t)__ffs(bal) - 1u : (uint32_t)KB_NT;
            __syncthreads();
            if (warp == 0) {
                const unsigned has = __ballot_sync(0xffffffffu, s_w[lane] != (uint32_t)KB_NT);
                if (lane == 0) s_first = has ? s_w[__ffs(has) - 1] : (uint32_t)KB_NT;
            }
            __syncthreads();
            const uint32_t first = s_first;
            if (first == (uint32_t)KB_NT) a = a + ((b - a - 1) / step) * step + 1; // every probe holds: M lies behind the last one
            else { b = a + first * step; if (first) a = a + (first - 1) * step + 1; }
            __syncthreads(); // s_w / s_first are rewritten by the next round
        }
        const uint32_t M = a;
        for (uint32_t m = tid; m < M; m += KB_NT) kd_swap_rec(rec, An[m], Bn[nB - 1 - m]);
        __syncthreads();
        if (tid == 0) {
            uint32_t istar;
            if (M < nA && (M == 0 || An[M] < Bn[nB - M])) istar = An[M];
            else if (M > 0) istar = Bn[nB - M];
            else istar = top;
            if (istar != top) kd_swap_rec(rec, istar, top);
            uint32_t nlo = lo, nhi = hi;
            if (istar >= k) nhi = istar;
            if (istar <= k) nlo = istar + 1;
            s_lo = nlo; s_hi = nhi;
        }
        __syncthreads();
        lo = s_lo; hi = s_hi;
    }
}

// One CTA arranges one node of more than KD_SMALL particles in global memory.  A, B: scratch of nL entries each (a node uses
// the entries of its own position range).  No __restrict__ / __ldg on rec: the CTA re-reads what it wrote.
__global__ void __launch_bounds__(KB_NT) k_kd_big(KdRec *rec, const uint32_t *__restrict__ nstart, const uint32_t *__restrict__ nend,
                                                  const uint32_t *__restrict__ list, uint32_t *A, uint32_t *B, uint32_t skip_from)
{
    __shared__ double s_red[6][KB_NT / 32];
    __shared__ int s_dim;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t nd = list[blockIdx.x];
    const uint32_t s = nstart[nd], e = nend[nd];
    if (e - s >= skip_from) return; // arranged by k_kd_huge (whole grid or cluster)
    // ---- split dimension: largest spread, the first dimension wins a tie (KDTree.cxx:463-503) ----
    {
        double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
        for (uint32_t p = s + tid; p < e; p += KB_NT) {
            const double2 a = *reinterpret_cast<const double2 *>(rec + p);
            const double c = rec[p].z;
            mn[0] = fmin(mn[0], a.x); mx[0] = fmax(mx[0], a.x);
            mn[1] = fmin(mn[1], a.y); mx[1] = fmax(mx[1], a.y);
            mn[2] = fmin(mn[2], c); mx[2] = fmax(mx[2], c);
        }
        for (int k = 0; k < 3; k++) {
            for (int o = 16; o; o >>= 1) {
                mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
                mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
            }
            if (lane == 0) { s_red[k][warp] = mn[k]; s_red[3 + k][warp] = mx[k]; }
        }
        __syncthreads();
        if (tid == 0) {
            double lo3[3], hi3[3];
            for (int k = 0; k < 3; k++) {
                lo3[k] = s_red[k][0]; hi3[k] = s_red[3 + k][0];
                for (int w = 1; w < KB_NT / 32; w++) { lo3[k] = fmin(lo3[k], s_red[k][w]); hi3[k] = fmax(hi3[k], s_red[3 + k][w]); }
            }
            int d = 0;
            double best = __dsub_rn(hi3[0], lo3[0]);
            for (int k = 1; k < 3; k++) {
                const double sp = __dsub_rn(hi3[k], lo3[k]);
                if (sp > best) { best = sp; d = k; }
            }
            s_dim = d;
        }
        __syncthreads();
    }
    const int d = s_dim;
    kd_cta_select(rec, s + (e - s - 1) / 2, d, s, e, A + s, B + s);
}

// ---- nodes of at least KD_HUGE particles: the whole GPU on one node ----------------------------------------------------------
// The same iteration as k_kd_big, with the scans of [lo, top) cut into one contiguous slice per CTA and a grid barrier between
// its phases (cooperative launch): count the two kinds of stops per slice, prefix over the slices, write the lists A / B, find M
// (every CTA for itself), swap the M pairs (all threads of the grid), place the pivot (one thread).  A single CTA moves ~0.2 G
// particles per second through its selections, so the root of a 5e7-member group alone took 235 ms of the 470 ms build; with
// 148 CTAs the top levels cost what their memory traffic costs.  Everything another CTA wrote is read with ld.cg (L1 is not
// coherent between SMs).
constexpr uint32_t KD_HUGE = 1u << 22;      // default: nodes of at least 4 Mi particles (VRGPU_KD_HUGE=<particles>, 0 = never)
constexpr uint32_t KD_HUGE_TAIL = 1u << 16; // ranges of at most this length are finished by one CTA
constexpr uint32_t KD_CLUSTER_MIN = 1u << 18, KD_CLUSTER_CTAS = 8; // the middle tier: a cluster of 8 CTAs per node of at least 256 Ki particles
struct KdHugeState {
    unsigned long long mn[3], mx[3];   // ordered images of the bounding box (pt_ord)
    unsigned long long cnt[256];       // per CTA: stops of kind A | kind B << 32 in its slice
    double x;
    uint32_t lo, hi;
};
__device__ __forceinline__ double kd_coord_cg(const KdRec *rec, uint32_t p, int d) { return __ldcg(reinterpret_cast<const double *>(rec + p) + d); }
__device__ __forceinline__ void kd_swap_rec_cg(KdRec *rec, uint32_t a, uint32_t b)
{
    uint4 *pa = reinterpret_cast<uint4 *>(rec + a), *pb = reinterpret_cast<uint4 *>(rec + b);
    const uint4 a0 = __ldcg(pa), a1 = __ldcg(pa + 1), b0 = __ldcg(pb), b1 = __ldcg(pb + 1);
    pa[0] = b0; pa[1] = b1; pb[0] = a0; pb[1] = a1;
}
// CLUSTER = false: the whole grid works through the nodes of the list one after the other (cooperative launch, grid barrier).
// CLUSTER = true: one thread-block cluster per list entry, all of them at once (hardware cluster barrier): the middle tier, nodes of
// at least min_size and less than max_size particles.
template <bool CLUSTER> struct KdSync;
template <> struct KdSync<false> {
    cooperative_groups::grid_group g = cooperative_groups::this_grid();
    __device__ void sync() { g.sync(); }
    __device__ uint32_t size() const { return gridDim.x; }
    __device__ uint32_t rank() const { return blockIdx.x; }
};
template <> struct KdSync<true> {
    cooperative_groups::cluster_group g = cooperative_groups::this_cluster();
    __device__ void sync() { g.sync(); }
    __device__ uint32_t size() const { return g.num_blocks(); }
    __device__ uint32_t rank() const { return g.block_rank(); }
};
template <bool CLUSTER>
__global__ void __launch_bounds__(KB_NT) k_kd_huge(KdRec *rec, const uint32_t *__restrict__ nstart, const uint32_t *__restrict__ nend,
                                                   const uint32_t *__restrict__ list, uint32_t nlist, uint32_t *A, uint32_t *B,
                                                   KdHugeState *st_all, uint32_t min_size, uint32_t max_size)
{
    KdSync<CLUSTER> grid;
    __shared__ double s_red[6][KB_NT / 32];
    __shared__ uint32_t s_w[KB_NT / 32];
    __shared__ uint32_t s_tot, s_first;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t G = grid.size(), c = grid.rank();
    KdHugeState *st = CLUSTER ? st_all + blockIdx.x / G : st_all;
    volatile KdHugeState *vst = st;
    const uint32_t li0 = CLUSTER ? blockIdx.x / G : 0u, li1 = CLUSTER ? li0 + 1u : nlist;
    for (uint32_t li = li0; li < li1; li++) {
        const uint32_t nd = list[li];
        const uint32_t s = nstart[nd], e = nend[nd];
        if (e - s < min_size || e - s >= max_size) continue; // (the same decision in every CTA of the grid / cluster)
        // ---- split dimension: largest spread, the first dimension wins a tie (KDTree.cxx:463-503) ----
        if (c == 0 && tid < 3) { st->mn[tid] = ~0ull; st->mx[tid] = 0ull; }
        grid.sync();
        {
            double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
            for (uint32_t p = s + c * KB_NT + tid; p < e; p += G * KB_NT) {
                const double a = kd_coord_cg(rec, p, 0), b = kd_coord_cg(rec, p, 1), cc = kd_coord_cg(rec, p, 2);
                mn[0] = fmin(mn[0], a); mx[0] = fmax(mx[0], a);
                mn[1] = fmin(mn[1], b); mx[1] = fmax(mx[1], b);
                mn[2] = fmin(mn[2], cc); mx[2] = fmax(mx[2], cc);
            }
            for (int k = 0; k < 3; k++) {
                for (int o = 16; o; o >>= 1) {
                    mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
                    mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
                }
                if (lane == 0) { s_red[k][warp] = mn[k]; s_red[3 + k][warp] = mx[k]; }
            }
            __syncthreads();
            if (tid < 3) {
                double lo1 = s_red[tid][0], hi1 = s_red[3 + tid][0];
                for (int w = 1; w < KB_NT / 32; w++) { lo1 = fmin(lo1, s_red[tid][w]); hi1 = fmax(hi1, s_red[3 + tid][w]); }
                atomicMin(&st->mn[tid], pt_ord(lo1));
                atomicMax(&st->mx[tid], pt_ord(hi1));
            }
        }
        grid.sync();
        int d = 0;
        {
            double best = __dsub_rn(pt_unord(vst->mx[0]), pt_unord(vst->mn[0]));
            for (int k = 1; k < 3; k++) {
                const double sp = __dsub_rn(pt_unord(vst->mx[k]), pt_unord(vst->mn[k]));
                if (sp > best) { best = sp; d = k; }
            }
        }
        const uint32_t k = s + (e - s - 1) / 2;
        uint32_t lo = s, hi = e;
        uint32_t *An = A + s, *Bn = B + s;
        // the pivot of an iteration is taken (and parked at the end of the range) by the thread that placed the previous one
        if (c == 0 && tid == 0) {
            st->x = kd_coord_cg(rec, k, d);
            if (k != hi - 1) kd_swap_rec_cg(rec, k, hi - 1);
        }
        grid.sync();
        while (lo + 1 < hi && hi - lo > KD_HUGE_TAIL) {
            const double x = vst->x;
            const uint32_t top = hi - 1; // the scans cover [lo, top)
            // this CTA's slice: whole tiles, contiguous, in the order of the CTA index
            const uint32_t ntiles = (top - lo + KB_TILE - 1) / KB_TILE;
            const uint32_t tper = (ntiles + G - 1) / G;
            const uint32_t t0 = min(ntiles, c * tper), t1 = min(ntiles, t0 + tper);
            const uint32_t my0 = lo + t0 * KB_TILE, my1 = min(top, lo + t1 * KB_TILE);
            // ---- phase 1: stops in my slice ----
            {
                uint32_t ca = 0, cb = 0;
                for (uint32_t p = my0 + tid; p < my1; p += KB_NT) {
                    const double v = kd_coord_cg(rec, p, d);
                    ca += v >= x ? 1u : 0u;
                    cb += v <= x ? 1u : 0u;
                }
                for (int o = 16; o; o >>= 1) { ca += __shfl_xor_sync(0xffffffffu, ca, o); cb += __shfl_xor_sync(0xffffffffu, cb, o); }
                if (lane == 0) { s_w[warp] = ca; s_red[0][warp] = (double)cb; }
                __syncthreads();
                if (tid == 0) {
                    unsigned long long ta = 0, tb = 0;
                    for (int w = 0; w < KB_NT / 32; w++) { ta += s_w[w]; tb += (unsigned long long)s_red[0][w]; }
                    st->cnt[c] = ta | (tb << 32);
                }
                __syncthreads();
            }
            grid.sync();
            // ---- prefix over the slices ----
            uint32_t nA = 0, nB = 0, baseA = 0, baseB = 0;
            for (uint32_t cc = 0; cc < G; cc++) {
                const unsigned long long v = vst->cnt[cc];
                if (cc == c) { baseA = nA; baseB = nB; }
                nA += (uint32_t)(v & 0xffffffffull); nB += (uint32_t)(v >> 32);
            }
            // ---- phase 2: the two lists (both ascending; B is read from its end) ----
            {
                uint32_t wa = baseA, wb = baseB;
                for (uint32_t base = my0; base < my1; base += KB_TILE) {
                    const uint32_t p0 = base + tid * KB_ITEMS;
                    unsigned fa = 0, fb = 0;
#pragma unroll
                    for (int j = 0; j < KB_ITEMS; j++)
                        if (p0 + j < my1) { const double v = kd_coord_cg(rec, p0 + j, d); fa |= (v >= x ? 1u : 0u) << j; fb |= (v <= x ? 1u : 0u) << j; }
                    const uint32_t mine = (uint32_t)__popc(fa) | ((uint32_t)__popc(fb) << 16);
                    uint32_t incl = mine;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= (unsigned)o) incl += y;
                    }
                    if (lane == 31) s_w[warp] = incl;
                    __syncthreads();
                    if (warp == 0) {
                        const uint32_t wv = s_w[lane];
                        uint32_t wi = wv;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const uint32_t y = __shfl_up_sync(0xffffffffu, wi, o);
                            if (lane >= (unsigned)o) wi += y;
                        }
                        s_w[lane] = wi - wv;
                        if (lane == 31) s_tot = wi;
                    }
                    __syncthreads();
                    const uint32_t excl = s_w[warp] + incl - mine, tot = s_tot;
                    uint32_t ea = wa + (excl & 0xffffu), eb = wb + (excl >> 16);
#pragma unroll
                    for (int j = 0; j < KB_ITEMS; j++) {
                        if ((fa >> j) & 1u) An[ea++] = p0 + j;
                        if ((fb >> j) & 1u) Bn[eb++] = p0 + j;
                    }
                    wa += tot & 0xffffu; wb += tot >> 16;
                    __syncthreads(); // s_w / s_tot are rewritten by the next tile
                }
            }
            grid.sync();
            // ---- M = first m with !(A[m] < B[nB-1-m]): every CTA searches for itself ----
            const uint32_t mm = nA < nB ? nA : nB;
            uint32_t a = 0, b = mm;
            while (a < b) {
                const uint32_t step = (b - a + KB_NT - 1) / KB_NT;
                const uint32_t m = a + tid * step;
                const bool fail = m < b && !(__ldcg(An + m) < __ldcg(Bn + (nB - 1 - m)));
                const unsigned bal = __ballot_sync(0xffffffffu, fail);
                if (lane == 0) s_w[warp] = bal ? warp * 32 + (uint32_t)__ffs(bal) - 1u : (uint32_t)KB_NT;
                __syncthreads();
                if (warp == 0) {
                    const unsigned has = __ballot_sync(0xffffffffu, s_w[lane] != (uint32_t)KB_NT);
                    if (lane == 0) s_first = has ? s_w[__ffs(has) - 1] : (uint32_t)KB_NT;
                }
                __syncthreads();
                const uint32_t first = s_first;
                if (first == (uint32_t)KB_NT) a = a + ((b - a - 1) / step) * step + 1;
                else { b = a + first * step; if (first) a = a + (first - 1) * step + 1; }
                __syncthreads();
            }
            const uint32_t M = a;
            for (uint32_t m = c * KB_NT + tid; m < M; m += G * KB_NT) kd_swap_rec_cg(rec, __ldcg(An + m), __ldcg(Bn + (nB - 1 - m)));
            grid.sync();
            if (c == 0 && tid == 0) {
                uint32_t istar;
                if (M < nA && (M == 0 || __ldcg(An + M) < __ldcg(Bn + (nB - M)))) istar = __ldcg(An + M);
                else if (M > 0) istar = __ldcg(Bn + (nB - M));
                else istar = top;
                if (istar != top) kd_swap_rec_cg(rec, istar, top);
                uint32_t nlo = lo, nhi = hi;
                if (istar >= k) nhi = istar;
                if (istar <= k) nlo = istar + 1;
                st->lo = nlo; st->hi = nhi;
                if (nlo + 1 < nhi && nhi - nlo > KD_HUGE_TAIL) { // the next cooperative iteration's pivot
                    st->x = kd_coord_cg(rec, k, d);
                    if (k != nhi - 1) kd_swap_rec_cg(rec, k, nhi - 1);
                }
            }
            grid.sync();
            lo = vst->lo; hi = vst->hi;
        }
        // the rest of the selection is too short to be worth a grid barrier per phase: one CTA finishes it (plain loads are safe: this
        // kernel has only ever read rec with ld.cg, so no stale line can sit in this SM's L1)
        if (c == 0 && lo + 1 < hi) kd_cta_select(rec, k, d, lo, hi, An, Bn);
        grid.sync(); // the next node's box reset must not overtake readers of this node's state; the tail's writes are visible
    }
}

// One CTA finishes one subtree of at most KD_SMALL particles in shared memory: columns + member (SoA, conflict-free scans),
// one warp per node, sub-level by sub-level.
constexpr int KS_NT = 256, KS_NW = KS_NT / 32, KS_LIST = KD_SMALL / 2;
struct KsShared {
    double x[KD_SMALL], y[KD_SMALL], z[KD_SMALL];
    uint32_t member[KD_SMALL];
    uint16_t A[KD_SMALL], B[KD_SMALL];
    uint16_t ls[2][KS_LIST], le[2][KS_LIST];
    uint32_t ncur, nnext;
};
__device__ __forceinline__ void ks_swap(KsShared &S, uint32_t a, uint32_t b)
{
    const double ax = S.x[a], ay = S.y[a], az = S.z[a];
    const uint32_t am = S.member[a];
    S.x[a] = S.x[b]; S.y[a] = S.y[b]; S.z[a] = S.z[b]; S.member[a] = S.member[b];
    S.x[b] = ax; S.y[b] = ay; S.z[b] = az; S.member[b] = am;
}
// MedianPos on the node [s, e) by one warp; all lanes hold the same lo / hi / counts
__device__ __forceinline__ void ks_warp_select(KsShared &S, const uint32_t s, const uint32_t e, const unsigned lane)
{
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    for (uint32_t p = s + lane; p < e; p += 32) {
        mn[0] = fmin(mn[0], S.x[p]); mx[0] = fmax(mx[0], S.x[p]);
        mn[1] = fmin(mn[1], S.y[p]); mx[1] = fmax(mx[1], S.y[p]);
        mn[2] = fmin(mn[2], S.z[p]); mx[2] = fmax(mx[2], S.z[p]);
    }
    int d = 0;
    double best = 0;
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o; o >>= 1) {
            mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
        const double sp = __dsub_rn(mx[k], mn[k]);
        if (k == 0 || sp > best) { best = sp; d = k; }
    }
    const double *val = d == 0 ? S.x : (d == 1 ? S.y : S.z);
    const uint32_t k = s + (e - s - 1) / 2;
    const unsigned lt = (1u << lane) - 1u;
    uint32_t lo = s, hi = e;
    uint16_t *An = S.A + s, *Bn = S.B + s;
    while (lo + 1 < hi) {
        const double x = val[k];
        const uint32_t top = hi - 1;
        __syncwarp();
        if (lane == 0 && k != top) ks_swap(S, k, top);
        __syncwarp();
        uint32_t nA = 0, nB = 0;
        for (uint32_t base = lo; base < top; base += 32) {
            const uint32_t p = base + lane;
            const bool in = p < top;
            const double v = in ? val[p] : 0.0;
            const bool ge = in && v >= x, le = in && v <= x;
            const unsigned ba = __ballot_sync(0xffffffffu, ge), bb = __ballot_sync(0xffffffffu, le);
            if (ge) An[nA + __popc(ba & lt)] = (uint16_t)p;
            if (le) Bn[nB + __popc(bb & lt)] = (uint16_t)p;
            nA += __popc(ba); nB += __popc(bb);
        }
        __syncwarp();
        const uint32_t mm = nA < nB ? nA : nB;
        uint32_t M = 0;
        for (uint32_t base = 0; base < mm; base += 32) {
            const uint32_t m = base + lane;
            uint32_t pa = 0, pb = 0;
            bool ok = false;
            if (m < mm) { pa = An[m]; pb = Bn[nB - 1 - m]; ok = pa < pb; }
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (ok) ks_swap(S, pa, pb); // the swapped positions are pairwise distinct
            const uint32_t c = (uint32_t)__popc(bal);
            M += c;
            if (c < 32) break;
        }
        __syncwarp();
        uint32_t istar;
        if (M < nA && (M == 0 || An[M] < Bn[nB - M])) istar = An[M];
        else if (M > 0) istar = Bn[nB - M];
        else istar = top;
        __syncwarp();
        if (lane == 0 && istar != top) ks_swap(S, istar, top);
        __syncwarp();
        if (istar >= k) hi = istar;
        if (istar <= k) lo = istar + 1;
    }
}
__global__ void __launch_bounds__(KS_NT) k_kd_small(KdRec *rec, const uint32_t *__restrict__ nstart, const uint32_t *__restrict__ nend,
                                                    const uint32_t *__restrict__ list, const uint32_t *__restrict__ count, uint32_t bucket)
{
    extern __shared__ __align__(16) unsigned char ks_raw[];
    KsShared &S = *reinterpret_cast<KsShared *>(ks_raw);
    if (blockIdx.x >= *count) return;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t nd = list[blockIdx.x];
    const uint32_t g0 = nstart[nd], n = nend[nd] - g0;
    for (uint32_t p = tid; p < n; p += KS_NT) {
        const uint4 *q = reinterpret_cast<const uint4 *>(rec + g0 + p);
        const uint4 a = q[0], b = q[1];
        S.x[p] = __hiloint2double((int)a.y, (int)a.x);
        S.y[p] = __hiloint2double((int)a.w, (int)a.z);
        S.z[p] = __hiloint2double((int)b.y, (int)b.x);
        S.member[p] = b.z;
    }
    if (tid == 0) { S.ls[0][0] = 0; S.le[0][0] = (uint16_t)n; S.ncur = n > bucket ? 1u : 0u; }
    int cur = 0;
    for (;;) {
        __syncthreads();
        const uint32_t ncur = S.ncur;
        if (ncur == 0) break;
        __syncthreads();
        if (tid == 0) S.nnext = 0;
        __syncthreads();
        for (uint32_t i = warp; i < ncur; i += KS_NW) {
            const uint32_t s = S.ls[cur][i], e = S.le[cur][i];
            ks_warp_select(S, s, e, lane);
            if (lane == 0) {
                const uint32_t mid = s + (e - s - 1) / 2 + 1;
                if (mid - s > bucket) { const uint32_t t = atomicAdd(&S.nnext, 1u); S.ls[cur ^ 1][t] = (uint16_t)s; S.le[cur ^ 1][t] = (uint16_t)mid; }
                if (e - mid > bucket) { const uint32_t t = atomicAdd(&S.nnext, 1u); S.ls[cur ^ 1][t] = (uint16_t)mid; S.le[cur ^ 1][t] = (uint16_t)e; }
            }
        }
        __syncthreads();
        if (tid == 0) S.ncur = S.nnext;
        cur ^= 1;
    }
    for (uint32_t p = tid; p < n; p += KS_NT) {
        KdRec r;
        r.x = S.x[p]; r.y = S.y[p]; r.z = S.z[p]; r.member = S.member[p]; r.pad = 0;
        rec[g0 + p] = r;
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
// Internal nodes: the reference sums EVERY node from its first particle to its last (unbind.cxx:1672-1675), not parent =
// left + right, and which cells a walk opens can depend on the last bit of the centre of mass (float32-quantised inputs
// produce exact ties in the opening test: 68 of 200 000 particles of tests/test_gpu_tree.py::test_tree_with_tied_coordinates
// flip with pairwise sums).  So every node gets the sequential sum:
//   nodes above KD_SMALL particles: one CTA per (node, component) with the exact integer scan of seqsum.h,
//   subtrees of at most KD_SMALL particles: one CTA per subtree, its particles in shared memory, one thread per (node, component).
__global__ void __launch_bounds__(SQ_NT) k_pt_big_sums(PtNodes N, const uint32_t *__restrict__ list, const double *__restrict__ tx,
                                                       const double *__restrict__ ty, const double *__restrict__ tz, const double *__restrict__ tm)
{
    __shared__ SeqShared sh;
    const uint32_t nd = list[blockIdx.x >> 2], comp = blockIdx.x & 3u;
    // The node's sum over its particles in order is its left child's sum carried on over the right child's particles: the same
    // sequence of roundings, half the elements.  The levels run bottom-up (one launch each), so the left child -- a big node of the
    // level below, or the root of a shared-memory subtree -- is final.
    const uint32_t lc = N.left[nd];
    const uint32_t s = N.start[lc + 1];
    const unsigned long long n = N.end[lc + 1] - s;
    const double s0 = (comp == 0 ? N.sx : (comp == 1 ? N.sy : (comp == 2 ? N.sz : N.sm)))[lc];
    double r;
    if (comp == 3) r = block_seqsum(TermMass{tm + s}, n, s0, sh);
    else r = block_seqsum(TermMomentum{(comp == 0 ? tx : (comp == 1 ? ty : tz)) + s, tm + s}, n, s0, sh);
    if (threadIdx.x == 0) (comp == 0 ? N.sx : (comp == 1 ? N.sy : (comp == 2 ? N.sz : N.sm)))[nd] = r;
}

constexpr int PS_NT = 256, PS_QUEUE = KD_SMALL; // a subtree of n particles has fewer than n internal nodes
struct PsShared {
    double x[KD_SMALL], y[KD_SMALL], z[KD_SMALL], m[KD_SMALL];
    uint32_t qnd[PS_QUEUE];
    uint16_t qs[PS_QUEUE], qe[PS_QUEUE];
    uint32_t qn;
};
__global__ void __launch_bounds__(PS_NT) k_pt_sub_sums(PtNodes N, const uint32_t *__restrict__ list, const uint32_t *__restrict__ count,
                                                       const double *__restrict__ tx, const double *__restrict__ ty, const double *__restrict__ tz,
                                                       const double *__restrict__ tm)
{
    extern __shared__ __align__(16) unsigned char ps_raw[];
    PsShared &S = *reinterpret_cast<PsShared *>(ps_raw);
    if (blockIdx.x >= *count) return;
    const unsigned tid = threadIdx.x;
    const uint32_t root = list[blockIdx.x];
    const uint32_t g0 = N.start[root], n = N.end[root] - g0;
    for (uint32_t p = tid; p < n; p += PS_NT) { S.x[p] = tx[g0 + p]; S.y[p] = ty[g0 + p]; S.z[p] = tz[g0 + p]; S.m[p] = tm[g0 + p]; }
    if (tid == 0) { S.qnd[0] = root; S.qs[0] = 0; S.qe[0] = (uint16_t)n; S.qn = 1; } // the root splits (it is on the list)
    __syncthreads();
    // internal nodes of the subtree, level by level (children of the nodes queued by the previous round)
    uint32_t qb = 0, qe = 1;
    while (qb < qe) {
        for (uint32_t i = qb + tid; i < qe; i += PS_NT) {
            const uint32_t l = N.left[S.qnd[i]];
            for (uint32_t c = l; c <= l + 1; c++)
                if (N.left[c] != NOCHILD) {
                    const uint32_t t = atomicAdd(&S.qn, 1u);
                    S.qnd[t] = c; S.qs[t] = (uint16_t)(N.start[c] - g0); S.qe[t] = (uint16_t)(N.end[c] - g0);
                }
        }
        __syncthreads();
        qb = qe;
        qe = S.qn;
        __syncthreads();
    }
    // one thread per (node, component); the queue is in breadth-first order, so the long sums are spread over the threads
    for (uint32_t task = tid; task < 4 * qe; task += PS_NT) {
        const uint32_t q = task >> 2, comp = task & 3u;
        const uint32_t a = S.qs[q], b = S.qe[q];
        double acc = 0.0;
        if (comp == 3) for (uint32_t p = a; p < b; p++) acc = __dadd_rn(acc, S.m[p]);
        else {
            const double *v = comp == 0 ? S.x : (comp == 1 ? S.y : S.z);
            for (uint32_t p = a; p < b; p++) acc = __dadd_rn(acc, __dmul_rn(v[p], S.m[p]));
        }
        (comp == 0 ? N.sx : (comp == 1 ? N.sy : (comp == 2 ? N.sz : N.sm)))[S.qnd[q]] = acc;
    }
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
// Measured and dropped (profiles/r02_walk.md): prefetch.global.L1 of both possible successors of a node -- 10 % slower; one
// warp per leaf with 8 targets x 4 sources and a shared-memory source stack -- the same number of warp steps (240 M against
// 251 M at 256^3), each dearer: 28.0 ms against 13.3 ms.
// STATS (VRGPU_WALK_STATS=1, a measuring aid): ninter[1..5] = warp-level node visits, visits in which some lane accepted the
// monopole, accepting lanes summed over those, warp-level leaf-particle iterations, opening lanes summed over those.
template <bool STATS>
__global__ void __launch_bounds__(128) k_pt_walk(const PtNode64 *__restrict__ nodes, const uint32_t *__restrict__ rootidx, uint32_t nL,
                                                 const double *__restrict__ tx, const double *__restrict__ ty, const double *__restrict__ tz,
                                                 const double *__restrict__ tm, double G, double eps2, const uint32_t *__restrict__ perm,
                                                 const uint32_t *__restrict__ slot_of, double *__restrict__ phi,
                                                 unsigned long long *__restrict__ ninter, const double2 *__restrict__ tpk /* (x, y), (z, m) per position */)
{
    unsigned long long st_visits = 0, st_accv = 0, st_accl = 0, st_leaf = 0, st_leafl = 0;
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = p < nL;
    const uint32_t myroot = valid ? rootidx[p] : NOCHILD;
    double x = 0, y = 0, z = 0, mj = 0;
    if (valid) { x = tx[p]; y = ty[p]; z = tz[p]; mj = tm[p]; }
    // acc gathers sum(M / r) without the target's own mass (one DFMA per term; the reference's mj * M / r differs from mj * acc in
    // the last bit of a term, far inside the 1e-10 the tests ask for); 32-bit count per lane (a lane meets < 2^32 sources)
    double acc = 0.0;
    unsigned int cnt = 0;
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
                acc = fma(-c.x, rsqrt_refined(r2 + eps2), acc);
                cnt++;
                wait = next;
            }
            if (STATS) {
                const unsigned am = __ballot_sync(0xffffffffu, act && !open);
                st_visits++; st_accv += am != 0; st_accl += __popc(am);
            }
            if (!__any_sync(0xffffffffu, open)) { cur = next; continue; }
            if (left != NOCHILD) { cur = left; continue; }
            if (STATS) { st_leaf += end - start; st_leafl += (unsigned long long)(end - start) * __popc(__ballot_sync(0xffffffffu, open)); }
            for (uint32_t l = start; l < end; l++) {
                const double2 la = __ldg(tpk + 2 * (size_t)l), lb = __ldg(tpk + 2 * (size_t)l + 1);
                if (open && l != p) {
                    const double ax = __dsub_rn(x, la.x), ay = __dsub_rn(y, la.y), az = __dsub_rn(z, lb.x);
                    const double q2 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(ax, ax), __dmul_rn(ay, ay)), __dmul_rn(az, az)), eps2);
                    acc = fma(-lb.y, rsqrt_refined(q2), acc);
                    cnt++;
                }
            }
            cur = next;
        }
    }
    if (valid) phi[slot_of[perm[p]]] = (acc * mj) * G;
    unsigned long long cntw = cnt;
    for (int o = 16; o; o >>= 1) cntw += __shfl_xor_sync(0xffffffffu, cntw, o);
    if ((threadIdx.x & 31) == 0 && cntw) atomicAdd(ninter, cntw);
    if (STATS && (threadIdx.x & 31) == 0) {
        atomicAdd(ninter + 1, st_visits); atomicAdd(ninter + 2, st_accv); atomicAdd(ninter + 3, st_accl);
        atomicAdd(ninter + 4, st_leaf); atomicAdd(ninter + 5, st_leafl);
    }
}

__global__ void k_pt_pack_targets(const double *__restrict__ tx, const double *__restrict__ ty, const double *__restrict__ tz, const double *__restrict__ tm,
                                  uint32_t nL, double2 *__restrict__ tpk)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nL) return;
    tpk[2 * (size_t)p] = make_double2(tx[p], ty[p]);
    tpk[2 * (size_t)p + 1] = make_double2(tz[p], tm[p]);
}

// tree-ordered columns: coordinates straight from the build's records, the mass through the member index
__global__ void k_pt_gather_tree(const KdRec *__restrict__ rec, const double *__restrict__ m, const uint32_t *__restrict__ slot_of, uint32_t nL,
                                 double *__restrict__ tx, double *__restrict__ ty, double *__restrict__ tz, double *__restrict__ tm)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nL) return;
    const KdRec r = rec[p];
    tx[p] = r.x; ty[p] = r.y; tz[p] = r.z; tm[p] = m[slot_of[r.member]];
}

} // namespace

// The build shared by the potentials (below) and the local velocity density (veldens.cu): members of the groups above
// pp_threshold in tree order (t.perm over compact member indices, t.slot_of to the group-major columns) and the nodes over
// them, numbered level by level (right child = left + 1).  False when no group qualifies.
bool pt_build(vrgpu_ctx *ctx, const GroupSet &gs, int bucket, long long pp_threshold, PtTree &t)
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
    flags.reset(); 

    // ---- the nodes: a function of the group sizes alone, laid down level by level ----
    const uint32_t minleaf = std::max(1u, ((uint32_t)bucket + 1u) / 2u);
    const size_t cap = 2 * ((size_t)nL / minleaf + 1) + nLG + 16;
    nstart.alloc(ctx, cap); nend.alloc(ctx, cap); nleft.alloc(ctx, cap);
    rootidx.alloc(ctx, nL);
    perm.alloc(ctx, nL);
    const size_t cap_list = 4 * ((size_t)nL / KD_SMALL + 1) + nLG + 64;
    DBuf<uint32_t> &big_list = t.big_list, &sub_list = t.sub_list, &counters = t.sub_count;
    big_list.alloc(ctx, cap_list); sub_list.alloc(ctx, cap_list); counters.alloc(ctx, 2);
    DBuf<uint32_t> lflags, lexcl;
    std::vector<uint32_t> big_off{0u}; // nodes above KD_SMALL of level l: big_list[big_off[l], big_off[l+1])
    uint32_t nsub = 0, nnodes = nLG;
    {
        KScope kb(ctx, "ptree.nodes");
        VR_LAUNCH(ctx, k_pt_roots, grid_for((size_t)NG + 1, T), T, 0, gs.off.p, gflag.p, gexcl.p, sexcl.p, NG, ntot, nstart.p, nend.p, nleft.p, perm.p /* unused frontier */);
        {   // tree position -> root node
            DBuf<uint32_t> heads(ctx, nL), hexcl(ctx, nL);
            VR_CUDA(cudaMemsetAsync(heads.p, 0, sizeof(uint32_t) * nL, ctx->stream));
            VR_LAUNCH(ctx, k_pt_iota, grid_for(nLG, T), T, 0, hexcl.p, nLG);
            VR_LAUNCH(ctx, k_pt_heads, grid_for(nLG, T), T, 0, hexcl.p, nLG, nstart.p, nL, heads.p);
            exclusive_scan_u32(ctx, heads.p, hexcl.p, nL, nullptr);
            VR_LAUNCH(ctx, k_pt_fidx, grid_for(nL, T), T, 0, heads.p, hexcl.p, nL, rootidx.p);
        }
        VR_CUDA(cudaMemsetAsync(counters.p, 0, 2 * sizeof(uint32_t), ctx->stream));
        level_off.assign({0u, nLG});
        for (int level = 0; level < 64; level++) {
            const uint32_t n0 = level_off[level], n1 = level_off[level + 1];
            if (n1 == n0) break;
            const uint32_t cnt = n1 - n0;
            lflags.alloc(ctx, cnt); lexcl.alloc(ctx, cnt);
            VR_LAUNCH(ctx, k_kd_level_flags, grid_for(cnt, T), T, 0, nstart.p, nend.p, n0, n1, (uint32_t)bucket, level == 0 ? 1 : 0, lflags.p,
                      big_list.p + big_off.back(), sub_list.p, counters.p);
            exclusive_scan_u32(ctx, lflags.p, lexcl.p, cnt, total.p);
            uint32_t h[3];
            VR_CUDA(cudaMemcpyAsync(h, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
            VR_CUDA(cudaMemcpyAsync(h + 1, counters.p, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
            const uint32_t nsplit = h[0];
            big_off.push_back(big_off.back() + h[1]);
            VR_REQUIRE(big_off.back() <= cap_list, VRGPU_ECUDA, "kd-tree: node list capacity exceeded");
            VR_CUDA(cudaMemsetAsync(counters.p, 0, sizeof(uint32_t), ctx->stream));
            if (nsplit == 0) { nsub = h[2]; break; }
            VR_REQUIRE((size_t)nnodes + 2 * (size_t)nsplit <= cap, VRGPU_ECUDA, "kd-tree node capacity exceeded");
            VR_LAUNCH(ctx, k_kd_level_children, grid_for(cnt, T), T, 0, n0, n1, lflags.p, lexcl.p, (uint32_t)bucket, nstart.p, nend.p, nleft.p,
                      sub_list.p, counters.p);
            nnodes += 2 * nsplit;
            level_off.push_back(nnodes);
            // subtree roots so far are read back with the next level's counters; the last level has no children above the bucket
        }
        if (nsub == 0) { // the loop ended through its level limit or an empty level: read the subtree count
            uint32_t h[2];
            VR_CUDA(cudaMemcpyAsync(h, counters.p, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
            nsub = h[1];
        }
        VR_REQUIRE(nsub <= cap_list, VRGPU_ECUDA, "kd-tree: subtree list capacity exceeded");
    }

    // ---- the arrangement: MedianPos on every node ----
    {
        KScope kb(ctx, "ptree.build");
        t.rec.alloc(ctx, (size_t)nL * 4); // 32-byte records
        KdRec *rec = reinterpret_cast<KdRec *>(t.rec.p);
        VR_LAUNCH(ctx, k_kd_init_rec, grid_for(nL, T), T, 0, gs.x.p, gs.y.p, gs.z.p, slot_of.p, nL, rec);
        if (big_off.back() > 0) {
            DBuf<uint32_t> A(ctx, nL), B(ctx, nL);
            // nodes of at least KD_HUGE particles take the whole GPU, one after the other (k_kd_huge, cooperative launch); a level can
            // only hold such a node while nL >> level reaches that size.  VRGPU_KD_HUGE=<particles> moves the threshold, 0 switches
            // the kernel off (every node to k_kd_big).
            uint32_t huge_min = KD_HUGE;
            if (const char *e = getenv("VRGPU_KD_HUGE")) huge_min = (uint32_t)std::max(0ll, atoll(e));
            if (huge_min == 0) huge_min = 0xffffffffu;
            huge_min = std::max(huge_min, 4u * KD_HUGE_TAIL);
            DBuf<KdHugeState> hst;
            int hgrid = 0;
            if (nL >= huge_min) {
                int per_sm = 0;
                VR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_kd_huge<false>, KB_NT, 0));
                hgrid = std::min(256, ctx->sm_count * std::max(per_sm, 0));
                if (hgrid >= 2) hst.alloc(ctx, 1);
            }
            // the middle tier: nodes from KD_CLUSTER_MIN particles up to the whole-GPU threshold get a cluster of 8 CTAs each, all nodes
            // of a level at once (VRGPU_KD_CLUSTER=<particles>, 0 = off)
            uint32_t cl_min = KD_CLUSTER_MIN;
            if (const char *e = getenv("VRGPU_KD_CLUSTER")) cl_min = (uint32_t)std::max(0ll, atoll(e));
            if (cl_min == 0) cl_min = 0xffffffffu;
            cl_min = std::max(cl_min, 4u * KD_HUGE_TAIL);
            const uint32_t cl_max = hgrid >= 2 ? huge_min : 0xffffffffu;
            DBuf<KdHugeState> cst;
            for (size_t l = 0; l + 1 < big_off.size(); l++) {
                const uint32_t nb = big_off[l + 1] - big_off[l];
                if (!nb) continue;
                const uint32_t level_max = nL >> std::min<size_t>(l, 31);   // no node of this level is larger
                const bool huge = hgrid >= 2 && level_max >= huge_min;
                const bool clus = level_max >= cl_min && cl_min < cl_max;
                if (huge) {
                    KdRec *a_rec = rec;
                    const uint32_t *a_ns = nstart.p, *a_ne = nend.p, *a_list = big_list.p + big_off[l];
                    uint32_t a_n = nb, *a_A = A.p, *a_B = B.p;
                    KdHugeState *a_st = hst.p;
                    uint32_t a_min = huge_min, a_max = 0xffffffffu;
                    void *args[] = {&a_rec, &a_ns, &a_ne, &a_list, &a_n, &a_A, &a_B, &a_st, &a_min, &a_max};
                    VR_CUDA(cudaLaunchCooperativeKernel((const void *)k_kd_huge<false>, dim3((unsigned)hgrid), dim3(KB_NT), args, 0, ctx->stream));
                    ctx->stats.kernel_launches++;
                }
                if (clus) {
                    if (cst.n < nb) cst.alloc(ctx, nb);
                    cudaLaunchConfig_t cfg = {};
                    cfg.gridDim = dim3(nb * KD_CLUSTER_CTAS);
                    cfg.blockDim = dim3(KB_NT);
                    cfg.stream = ctx->stream;
                    cudaLaunchAttribute at[1];
                    at[0].id = cudaLaunchAttributeClusterDimension;
                    at[0].val.clusterDim.x = KD_CLUSTER_CTAS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                    cfg.attrs = at;
                    cfg.numAttrs = 1;
                    VR_CUDA(cudaLaunchKernelEx(&cfg, k_kd_huge<true>, rec, (const uint32_t *)nstart.p, (const uint32_t *)nend.p,
                                               (const uint32_t *)(big_list.p + big_off[l]), nb, A.p, B.p, cst.p, cl_min, cl_max));
                    ctx->stats.kernel_launches++;
                }
                const uint32_t skip_from = clus ? cl_min : (huge ? huge_min : 0xffffffffu);
                VR_LAUNCH(ctx, k_kd_big, nb, KB_NT, 0, rec, nstart.p, nend.p, big_list.p + big_off[l], A.p, B.p, skip_from);
            }
        }
        if (nsub) {
            VR_CUDA(cudaFuncSetAttribute(k_kd_small, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(KsShared)));
            VR_LAUNCH(ctx, k_kd_small, nsub, KS_NT, sizeof(KsShared), rec, nstart.p, nend.p, sub_list.p, counters.p + 1, (uint32_t)bucket);
        }
        VR_LAUNCH(ctx, k_kd_perm, grid_for(nL, T), T, 0, rec, nL, perm.p);
    }
    t.nL = nL;
    t.nLG = nLG;
    t.nnodes = nnodes;
    t.nbig = big_off.back();
    t.big_off = big_off;
    t.nsub = nsub;
    return true;
}

void potential_tree(vrgpu_ctx *ctx, const GroupSet &gs, double G, double eps, double theta, int bucket, long long pp_threshold, double *phi)
{
    if (gs.ntot == 0 || gs.ngroups == 0 || pp_threshold < 0) return;
    VR_REQUIRE(theta > 0, VRGPU_EINVAL, "TreeThetaOpen must be positive");
    PtTree t;
    if (!pt_build(ctx, gs, bucket, pp_threshold, t)) return;
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
        VR_LAUNCH(ctx, k_pt_gather_tree, grid_for(nL, T), T, 0, reinterpret_cast<const KdRec *>(t.rec.p), gs.m.p, slot_of.p, nL, tx.p, ty.p, tz.p, tm.p);
        t.rec.reset();
        VR_LAUNCH(ctx, k_pt_leaf_sums, grid_for(nnodes, T), T, 0, N, nnodes, tx.p, ty.p, tz.p, tm.p);
        if (t.nsub) {
            VR_CUDA(cudaFuncSetAttribute(k_pt_sub_sums, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PsShared)));
            VR_LAUNCH(ctx, k_pt_sub_sums, t.nsub, PS_NT, sizeof(PsShared), N, t.sub_list.p, t.sub_count.p + 1, tx.p, ty.p, tz.p, tm.p);
        }
        for (size_t l = t.big_off.size() - 1; l-- > 0;) { // the nodes above 2048 particles, deepest level first
            const uint32_t nb = t.big_off[l + 1] - t.big_off[l];
            if (nb) VR_LAUNCH(ctx, k_pt_big_sums, 4u * nb, SQ_NT, 0, N, t.big_list.p + t.big_off[l], tx.p, ty.p, tz.p, tm.p);
        }
        VR_LAUNCH(ctx, k_pt_cm, grid_for(nnodes, T), T, 0, N, nnodes);
        VR_LAUNCH(ctx, k_pt_bmax, grid_for(nL, T), T, 0, N, rootidx.p, nL, tx.p, ty.p, tz.p);
        VR_LAUNCH(ctx, k_pt_r2max, grid_for(nnodes, T), T, 0, N, nnodes, theta);
    }
    // ---- walk ----
    DBuf<unsigned long long> ninter(ctx, 8);
    DBuf<PtNode64> packed(ctx, nnodes);
    VR_CUDA(cudaMemsetAsync(ninter.p, 0, 8 * sizeof(unsigned long long), ctx->stream));
    static const bool walk_stats = [] { const char *e = getenv("VRGPU_WALK_STATS"); return e && atoi(e) == 1; }();
    {
        KScope k(ctx, "ptree.walk");
        DBuf<uint32_t> next(ctx, nnodes);
        VR_CUDA(cudaMemsetAsync(next.p, 0xff, sizeof(uint32_t) * nLG, ctx->stream)); // roots: nothing follows
        for (size_t l = 0; l + 1 < level_off.size(); l++) {
            const uint32_t n0 = level_off[l], n1 = level_off[l + 1];
            if (n1 > n0) VR_LAUNCH(ctx, k_pt_next, grid_for(n1 - n0, T), T, 0, nleft.p, next.p, n0, n1);
        }
        VR_LAUNCH(ctx, k_pt_pack, grid_for(nnodes, T), T, 0, N, next.p, nnodes, packed.p);
        DBuf<double2> tpk(ctx, 2 * (size_t)nL);
        VR_LAUNCH(ctx, k_pt_pack_targets, grid_for(nL, T), T, 0, tx.p, ty.p, tz.p, tm.p, nL, tpk.p);
        if (walk_stats)
            VR_LAUNCH(ctx, k_pt_walk<true>, grid_for(nL, 128), 128, 0, packed.p, rootidx.p, nL, tx.p, ty.p, tz.p, tm.p, G, eps * eps, perm.p, slot_of.p,
                      phi, ninter.p, tpk.p);
        else
            VR_LAUNCH(ctx, k_pt_walk<false>, grid_for(nL, 128), 128, 0, packed.p, rootidx.p, nL, tx.p, ty.p, tz.p, tm.p, G, eps * eps, perm.p, slot_of.p,
                      phi, ninter.p, tpk.p);
    }
    unsigned long long hs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    VR_CUDA(cudaMemcpyAsync(hs, ninter.p, sizeof(hs), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    const unsigned long long hi = hs[0];
    if (walk_stats)
        fprintf(stderr, "[walk stats] targets %u nodes %u interactions %llu | warp visits %llu, with an accepting lane %llu (lanes %llu) | warp leaf iterations %llu (lanes %llu)\n",
                nL, nnodes, hs[0], hs[1], hs[2], hs[3], hs[4], hs[5]);
    ctx->stats.n_tree_interactions += (long long)hi;
    ctx->stats.n_tree_nodes += (long long)nnodes;
}

namespace {
__global__ void k_pt_order_out(const uint32_t *__restrict__ perm, const uint32_t *__restrict__ slot_of, const uint32_t *__restrict__ gid,
                               const uint32_t *__restrict__ off, uint32_t nL, long long *__restrict__ out)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nL) return;
    const uint32_t slot = slot_of[perm[p]];
    out[slot_of[p]] = (long long)(slot - off[gid[slot]]); // trees cover their groups' slots in order: position p <-> slot_of[p]
}
} // namespace

void tree_order_device(vrgpu_ctx *ctx, const GroupSet &gs, int bucket, long long *out_dev)
{
    PtTree t;
    if (!pt_build(ctx, gs, bucket, 0, t)) return;
    VR_REQUIRE(t.nL == gs.ntot, VRGPU_ECUDA, "tree order: empty groups are not allowed");
    VR_LAUNCH(ctx, k_pt_order_out, grid_for(t.nL, 256), 256, 0, t.perm.p, t.slot_of.p, gs.gid.p, gs.off.p, t.nL, out_dev);
}

} // namespace vrgpu
