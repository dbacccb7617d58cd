// velociraptor-stf_b200/csrc/fof6d.cu -- 6D phase-space FOF inside every 3D group.
//
// Replaces the 6DFOF block of SearchFullSet (src/search.cxx:400-698): regroup by 3D group (qsort by PID :434),
// per-group velocity dispersion (:472-499, or the single-scale variant :443-470 with its quirk Q1), ScalePhase
// (:565-568), a 6-D KDTree + KDTree::FOF(1.0) per group (:570-571), un-scaling (:573-575), id offsets (:659-669) and
// ReorderGroupIDs (:681-689).
//
// Device design (not the reference's): all 3D groups are processed in ONE set of launches.
//   1. group-major SoA (groups.cu), members in ascending particle index like the reference's stable qsort;
//   2. per-group dispersion: one warp per group; the reference's summation ORDER is kept (a sequential chain in member
//      order, products rounded before the add) so the scale factors -- and with them every scaled coordinate -- are
//      bit-identical to the reference;
//   3. coordinates scaled in place (one rounding per coordinate), un-scaled copies written back (Q3);
//   4. cells of edge 1 in SCALED position space, key = (group, cz, cy, cx); radix sort + hash table;
//   5. lock-free union-find over PARTICLES: a cheap seeding pass (each particle against the head of its own cell)
//      builds the big components, a per-cell "uniform root" flag then lets the exhaustive pass skip whole cell pairs
//      that are already one component; the exhaustive pass (one warp per 32 particles, neighbour tiles staged in
//      shared memory) tests every remaining pair whose roots differ with the reference predicate
//         d2 = (dx^2+dy^2+dz^2) + (dvx^2+dvy^2+dvz^2) < 1   (KDLeafNode.cxx:556-558; fp64, left to right, no FMA);
//   6. component sizes / min pid -> ids by (size desc, min pid asc) over ALL groups (== per-group ids + offsets +
//      ReorderGroupIDs, canonicalised).
#include "groups.cuh"
#include "stages.cuh"
#include "seqsum.h"
#include "seqblock.cuh"
#include "ufhash.cuh"

#include <algorithm>
#include <memory>
#include <thread>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <cmath>

namespace vrgpu {

constexpr uint32_t NONE32 = 0xffffffffu;
constexpr int CBITS = 14;               // bits per scaled-cell coordinate in the key
constexpr uint64_t CMASK = (1ull << CBITS) - 1;

// ---------------------------------------------------------------------------------------------------
// per-group velocity scale
// ---------------------------------------------------------------------------------------------------
// Sequential-order sums.  The reference accumulates left to right (search.cxx:443-493) and 6DFOF membership depends on
// every bit of the dispersion, so the order is kept.  Groups below SQ_MIN members: one warp per group running the
// dependent DADD chain.  Larger groups: one CTA per group evaluating the same chain with the exact integer scan of
// seqsum.h (windows of SQ_W terms; plain adds only where the running sum changes binade).
// ---------------------------------------------------------------------------------------------------
struct TermDispersion { // pow(v_c - vmean_c, 2) * m for c = x, y, z of member k/3 (search.cxx:489)
    const double *vx, *vy, *vz, *m;
    double mx, my, mz;
    __device__ double operator()(unsigned long long k) const
    {
        const unsigned long long i = k / 3ull;
        const unsigned c = (unsigned)(k - 3ull * i);
        const double d = __dsub_rn(c == 0 ? vx[i] : (c == 1 ? vy[i] : vz[i]), c == 0 ? mx : (c == 1 ? my : mz));
        return __dmul_rn(__dmul_rn(d, d), m[i]);
    }
};

// ---------------------------------------------------------------------------------------------------
// Many-window evaluation of one long sequential sum (seqsum.h, "Many windows at once"): EXPERIMENT for the next measuring
// session, used only when VRGPU_DISP_MULTI_MIN names a member count (fof6d_device).  block_seqsum walks the windows of a
// group one after the other inside one CTA -- 295 ms for the 1.8 10^7-member group of the zoom configuration.  Here
//   k_seq_approx  one CTA per window: ordinary (inexact) sum of its terms;
//   k_seq_prefix  running sum at every window start (inexact; only its exponent is used: the predicted unit);
//   k_seq_spec    one CTA per window: the window's summary under the predicted unit (block scan of the term maps);
//   k_seq_walk    one CTA: thread 0 steps over the summaries (512 staged in shared memory at a time); a window whose
//                 summary does not apply is evaluated by the whole CTA with block_seqsum, exactly as before.
// The result is the sequential sum bit for bit whatever the predictions are (tests/host/seqsum_test.cpp runs the same
// logic with skewed and corrupted predictions).
// ---------------------------------------------------------------------------------------------------
static_assert(SQ_W <= (int)SEQ_SPEC_MAXTERMS, "a summarised window must not exceed SEQ_SPEC_MAXTERMS terms");

template <class Term> struct TermShift {
    Term t;
    unsigned long long base;
    __device__ double operator()(unsigned long long k) const { return t(base + k); }
};
struct TermMomentumD { // as TermMomentum
    const double *v, *m;
    __device__ double operator()(unsigned long long k) const { return __dmul_rn(v[k], m[k]); }
};
struct TermDispersionD { // as TermDispersion, means read from device memory (they are sequential sums themselves)
    const double *vx, *vy, *vz, *m;
    const double *mean; // 3
    __device__ double operator()(unsigned long long k) const
    {
        const unsigned long long i = k / 3ull;
        const unsigned c = (unsigned)(k - 3ull * i);
        const double d = __dsub_rn(c == 0 ? vx[i] : (c == 1 ? vy[i] : vz[i]), mean[c]);
        return __dmul_rn(__dmul_rn(d, d), m[i]);
    }
};

template <class Term> __global__ void __launch_bounds__(SQ_NT) k_seq_approx(Term term, unsigned long long n, double *__restrict__ wsum)
{
    __shared__ double part[SQ_NT / 32];
    const unsigned long long first = (unsigned long long)blockIdx.x * SQ_W;
    const unsigned w = (unsigned)min((unsigned long long)SQ_W, n - first);
    double a = 0.0;
    for (unsigned k = threadIdx.x; k < w; k += SQ_NT) a += term(first + k);
    for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int q = 0; q < SQ_NT / 32; q++) t += part[q];
        wsum[blockIdx.x] = t;
    }
}

__global__ void k_seq_prefix(const double *__restrict__ wsum, unsigned nwin, const double *__restrict__ s0, double *__restrict__ start)
{
    if (blockIdx.x || threadIdx.x) return;
    double run = s0 ? *s0 : 0.0;
    for (unsigned k = 0; k < nwin; k++) { start[k] = run; run += wsum[k]; }
}

template <class Term> __global__ void __launch_bounds__(SQ_NT) k_seq_spec(Term term, unsigned long long n, const double *__restrict__ start,
                                                                          SeqSpec *__restrict__ spec)
{
    __shared__ SeqMap wtot[SQ_NT / 32];
    __shared__ long long red[4][SQ_NT / 32];
    __shared__ int s_small;
    const unsigned tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    const unsigned long long first = (unsigned long long)blockIdx.x * SQ_W;
    const unsigned w = (unsigned)min((unsigned long long)SQ_W, n - first);
    const int eu = seq_predict_unit(start[blockIdx.x]);
    if (eu == SEQ_NOSPEC) { // uniform over the CTA
        if (tid == 0) { SeqSpec z{}; z.eu = SEQ_NOSPEC; spec[blockIdx.x] = z; }
        return;
    }
    if (tid == 0) s_small = 1;
    __syncthreads();
    const unsigned k0 = tid * SQ_PER;
    SeqMap tm[SQ_PER];
    SeqMap inc = seq_identity();
    bool small = true;
#pragma unroll
    for (int j = 0; j < SQ_PER; j++) {
        tm[j] = (k0 + j < w) ? seq_term_map(term(first + k0 + j), eu) : seq_identity();
        small = small && seq_term_small(tm[j]);
        inc = seq_compose(inc, tm[j]);
    }
    if (!small) s_small = 0; // benign race: every writer stores 0
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        SeqMap o;
        o.e = __shfl_up_sync(0xffffffffu, inc.e, d);
        o.o = __shfl_up_sync(0xffffffffu, inc.o, d);
        if ((int)lane >= d) inc = seq_compose(o, inc);
    }
    if (lane == 31) wtot[wid] = inc;
    __syncthreads();
    if (!s_small) { // a term that is not small against the predicted sum: no summary (additions could wrap)
        if (tid == 0) { SeqSpec z{}; z.eu = SEQ_NOSPEC; spec[blockIdx.x] = z; }
        return;
    }
    SeqMap pre = seq_identity();
    for (unsigned q = 0; q < wid; q++) pre = seq_compose(pre, wtot[q]);
    SeqMap ex;
    ex.e = __shfl_up_sync(0xffffffffu, inc.e, 1);
    ex.o = __shfl_up_sync(0xffffffffu, inc.o, 1);
    if (lane == 0) ex = seq_identity();
    pre = seq_compose(pre, ex);
    // extreme prefix increments of this thread's terms
    long long mine = SEQ_HI, maxe = -SEQ_HI, mino = SEQ_HI, maxo = -SEQ_HI;
#pragma unroll
    for (int j = 0; j < SQ_PER; j++) {
        if (k0 + j < w) {
            pre = seq_compose(pre, tm[j]);
            mine = min(mine, pre.e); maxe = max(maxe, pre.e);
            mino = min(mino, pre.o); maxo = max(maxo, pre.o);
        }
    }
    for (int o = 16; o; o >>= 1) {
        mine = min(mine, __shfl_xor_sync(0xffffffffu, mine, o)); maxe = max(maxe, __shfl_xor_sync(0xffffffffu, maxe, o));
        mino = min(mino, __shfl_xor_sync(0xffffffffu, mino, o)); maxo = max(maxo, __shfl_xor_sync(0xffffffffu, maxo, o));
    }
    if (lane == 0) { red[0][wid] = mine; red[1][wid] = maxe; red[2][wid] = mino; red[3][wid] = maxo; }
    __syncthreads();
    if (tid == 0) {
        SeqSpec r{};
        r.total = seq_identity();
        for (int q = 0; q < SQ_NT / 32; q++) r.total = seq_compose(r.total, wtot[q]);
        r.mine = red[0][0]; r.maxe = red[1][0]; r.mino = red[2][0]; r.maxo = red[3][0];
        for (int q = 1; q < SQ_NT / 32; q++) {
            r.mine = min(r.mine, red[0][q]); r.maxe = max(r.maxe, red[1][q]);
            r.mino = min(r.mino, red[2][q]); r.maxo = max(r.maxo, red[3][q]);
        }
        r.eu = eu;
        spec[blockIdx.x] = r;
    }
}

template <class Term> __global__ void __launch_bounds__(SQ_NT) k_seq_walk(Term term, unsigned long long n, const double *__restrict__ s0,
                                                                          const SeqSpec *__restrict__ spec, unsigned nwin, double *__restrict__ out,
                                                                          unsigned int *__restrict__ nslow)
{
    __shared__ SeqShared sh;
    __shared__ SeqSpec buf[SQ_NT];
    __shared__ unsigned s_taken;
    const unsigned tid = threadIdx.x;
    double s = s0 ? *s0 : 0.0;
    unsigned k = 0, slow = 0;
    while (k < nwin) {
        const unsigned cnt = min((unsigned)SQ_NT, nwin - k);
        if (tid < cnt) buf[tid] = spec[k + tid];
        __syncthreads();
        if (tid == 0) {
            unsigned j = 0;
            long long S;
            int eu;
            if (seq_decode_sum(s, S, eu)) {
                while (j < cnt && seq_spec_step(buf[j], S, eu)) j++;
                if (j) s = seq_encode_sum(S, eu);
            }
            sh.s = s;
            s_taken = j;
        }
        __syncthreads();
        s = sh.s;
        const unsigned j = s_taken;
        __syncthreads(); // everybody has read sh.s / s_taken before block_seqsum reuses sh
        k += j;
        if (j < cnt) { // the summary of window k does not apply: the exact windowed evaluation
            const unsigned long long first = (unsigned long long)k * SQ_W;
            s = block_seqsum(TermShift<Term>{term, first}, min((unsigned long long)SQ_W, n - first), s, sh);
            k++;
            slow++;
        }
    }
    if (tid == 0) {
        *out = s;
        if (nslow) atomicAdd(nslow, slow);
    }
}

// host side: enqueue the four launches; *out (device) receives fl(...fl(fl(*s0 + t0) + t1)...), s0 == nullptr means 0
struct SeqScratch {
    DBuf<double> wsum, start;
    DBuf<SeqSpec> spec;
    DBuf<unsigned int> nslow;
    unsigned cap = 0;
};
template <class Term> static void seq_multi_sum(vrgpu_ctx *ctx, SeqScratch &sc, Term term, unsigned long long n, const double *s0_dev, double *out_dev)
{
    const unsigned nwin = (unsigned)((n + SQ_W - 1) / SQ_W);
    VR_REQUIRE(nwin >= 1 && nwin <= sc.cap, VRGPU_ECUDA, "seq_multi_sum: scratch too small");
    VR_LAUNCH(ctx, k_seq_approx<Term>, nwin, SQ_NT, 0, term, n, sc.wsum.p);
    VR_LAUNCH(ctx, k_seq_prefix, 1, 32, 0, sc.wsum.p, nwin, s0_dev, sc.start.p);
    VR_LAUNCH(ctx, k_seq_spec<Term>, nwin, SQ_NT, 0, term, n, sc.start.p, sc.spec.p);
    VR_LAUNCH(ctx, k_seq_walk<Term>, 1, SQ_NT, 0, term, n, s0_dev, sc.spec.p, nwin, out_dev, sc.nslow.p);
}

// r = {sum m vx, sum m vy, sum m vz, sum m}  ->  mean[c] = r[c] / norm   (norm = r[3], or the single-scale mnorm of quirk Q1)
__global__ void k_seq_means(const double *__restrict__ r, double mnorm, int use_mnorm, double *__restrict__ mean)
{
    if (blockIdx.x || threadIdx.x) return;
    const double d = use_mnorm ? mnorm : r[3];
    mean[0] = __ddiv_rn(r[0], d); mean[1] = __ddiv_rn(r[1], d); mean[2] = __ddiv_rn(r[2], d);
}
// adaptive: vs2[g] = acc / sm * vfac * vfac (search.cxx:489-493); single scale: acc / mnorm (if > 0) * (vfac * vfac) for every chunk
__global__ void k_seq_finish(const double *__restrict__ r, double vfac, double mnorm, int single, uint32_t g_first, uint32_t g_count,
                             double *__restrict__ vs2)
{
    double acc = r[4];
    if (!single) acc = __dmul_rn(__dmul_rn(__ddiv_rn(acc, r[3]), vfac), vfac);
    else {
        if (mnorm > 0) acc = __ddiv_rn(acc, mnorm);
        acc = __dmul_rn(acc, __dmul_rn(vfac, vfac));
    }
    for (uint32_t g = g_first + blockIdx.x * blockDim.x + threadIdx.x; g < g_first + g_count; g += gridDim.x * blockDim.x) vs2[g] = acc;
}

// the dispersion of ONE member range [vx, vx + n) ... through the many-window path; vs2[g_first .. g_first + g_count) receive it
static void dispersion_multi(vrgpu_ctx *ctx, SeqScratch &sc, DBuf<double> &r /* 8 */, const double *vx, const double *vy, const double *vz,
                             const double *m, unsigned long long n, double vfac, bool single, double mnorm, uint32_t g_first, uint32_t g_count,
                             double *vs2)
{
    seq_multi_sum(ctx, sc, TermMomentumD{vx, m}, n, nullptr, r.p + 0);
    seq_multi_sum(ctx, sc, TermMomentumD{vy, m}, n, nullptr, r.p + 1);
    seq_multi_sum(ctx, sc, TermMomentumD{vz, m}, n, nullptr, r.p + 2);
    if (!single) seq_multi_sum(ctx, sc, TermMass{m}, n, nullptr, r.p + 3);
    VR_LAUNCH(ctx, k_seq_means, 1, 32, 0, r.p, mnorm, single ? 1 : 0, r.p + 5);
    seq_multi_sum(ctx, sc, TermDispersionD{vx, vy, vz, m, r.p + 5}, 3ull * n, nullptr, r.p + 4);
    VR_LAUNCH(ctx, k_seq_finish, 1, 256, 0, r.p, vfac, mnorm, single ? 1 : 0, g_first, g_count, vs2);
}

// FOF6DADAPTIVE (search.cxx:480-493): vmean = sum(m v)/sum(m); vscale2 = sum_k pow(v_k - vmean_k, 2) m / sum(m) * vfac^2.
// One CTA per group of at least SQ_MIN members (group ids are size-descending: the other CTAs leave at once).
__global__ void __launch_bounds__(SQ_NT) k_group_dispersion_big(const double *__restrict__ vx, const double *__restrict__ vy,
                                                                const double *__restrict__ vz, const double *__restrict__ m,
                                                                const uint32_t *__restrict__ off, double vfac, double *__restrict__ vs2,
                                                                uint32_t nskip /* leading groups done by the many-window path */)
{
    __shared__ SeqShared sh;
    const uint32_t g = blockIdx.x + 1;
    const uint32_t s0 = off[g], n = off[g + 1] - s0;
    if (n < SQ_MIN || g <= nskip) return;
    const double sx = block_seqsum(TermMomentum{vx + s0, m + s0}, n, 0.0, sh);
    const double sy = block_seqsum(TermMomentum{vy + s0, m + s0}, n, 0.0, sh);
    const double sz = block_seqsum(TermMomentum{vz + s0, m + s0}, n, 0.0, sh);
    const double sm = block_seqsum(TermMass{m + s0}, n, 0.0, sh);
    const TermDispersion td{vx + s0, vy + s0, vz + s0, m + s0, __ddiv_rn(sx, sm), __ddiv_rn(sy, sm), __ddiv_rn(sz, sm)};
    const double acc = block_seqsum(td, 3ull * n, 0.0, sh);
    if (threadIdx.x == 0) vs2[g] = __dmul_rn(__dmul_rn(__ddiv_rn(acc, sm), vfac), vfac);
}

// Groups below SQ_MIN members.  One warp per group; all lanes carry the same sequential sums.
__global__ void __launch_bounds__(256) k_group_dispersion(const double *__restrict__ vx, const double *__restrict__ vy,
                                                          const double *__restrict__ vz, const double *__restrict__ m,
                                                          const uint32_t *__restrict__ off, uint32_t ngroups, double vfac,
                                                          double *__restrict__ vs2)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t g = 1 + wg; g <= ngroups; g += nw) {
        const uint32_t s0 = off[g], s1 = off[g + 1];
        if (s1 - s0 >= SQ_MIN) continue; // k_group_dispersion_big
        double sx = 0, sy = 0, sz = 0, sm = 0;
        for (uint32_t base = s0; base < s1; base += 32) {
            const uint32_t s = base + lane;
            double px = 0, py = 0, pz = 0, mm = 0;
            if (s < s1) {
                mm = m[s];
                px = __dmul_rn(vx[s], mm); py = __dmul_rn(vy[s], mm); pz = __dmul_rn(vz[s], mm);
            }
            const int cnt = (int)min(32u, s1 - base);
            for (int t = 0; t < cnt; t++) {
                sx = __dadd_rn(sx, __shfl_sync(0xffffffffu, px, t));
                sy = __dadd_rn(sy, __shfl_sync(0xffffffffu, py, t));
                sz = __dadd_rn(sz, __shfl_sync(0xffffffffu, pz, t));
                sm = __dadd_rn(sm, __shfl_sync(0xffffffffu, mm, t));
            }
        }
        const double mx = __ddiv_rn(sx, sm), my = __ddiv_rn(sy, sm), mz = __ddiv_rn(sz, sm);
        double acc = 0;
        for (uint32_t base = s0; base < s1; base += 32) {
            const uint32_t s = base + lane;
            double tx = 0, ty = 0, tz = 0;
            if (s < s1) {
                const double mm = m[s];
                double d = __dsub_rn(vx[s], mx); tx = __dmul_rn(__dmul_rn(d, d), mm);
                d = __dsub_rn(vy[s], my); ty = __dmul_rn(__dmul_rn(d, d), mm);
                d = __dsub_rn(vz[s], mz); tz = __dmul_rn(__dmul_rn(d, d), mm);
            }
            const int cnt = (int)min(32u, s1 - base);
            for (int t = 0; t < cnt; t++) {
                acc = __dadd_rn(acc, __shfl_sync(0xffffffffu, tx, t));
                acc = __dadd_rn(acc, __shfl_sync(0xffffffffu, ty, t));
                acc = __dadd_rn(acc, __shfl_sync(0xffffffffu, tz, t));
            }
        }
        if (lane == 0) vs2[g] = __dmul_rn(__dmul_rn(__ddiv_rn(acc, sm), vfac), vfac);
    }
}

// FOF6D single scale (search.cxx:443-470) INCLUDING quirk Q1: the mass normalisation is ONE particle's mass (the
// particle that follows the largest group in the group-sorted array).  One CTA; result broadcast to every chunk.
__global__ void __launch_bounds__(SQ_NT) k_single_dispersion(const double *__restrict__ vx, const double *__restrict__ vy,
                                                             const double *__restrict__ vz, const double *__restrict__ m, uint32_t n1,
                                                             double mnorm, double vfac, uint32_t nchunks, double *__restrict__ vs2)
{
    __shared__ SeqShared sh;
    const double sx = block_seqsum(TermMomentum{vx, m}, n1, 0.0, sh);
    const double sy = block_seqsum(TermMomentum{vy, m}, n1, 0.0, sh);
    const double sz = block_seqsum(TermMomentum{vz, m}, n1, 0.0, sh);
    const TermDispersion td{vx, vy, vz, m, __ddiv_rn(sx, mnorm), __ddiv_rn(sy, mnorm), __ddiv_rn(sz, mnorm)};
    double acc = block_seqsum(td, 3ull * n1, 0.0, sh);
    if (mnorm > 0) acc = __ddiv_rn(acc, mnorm);
    acc = __dmul_rn(acc, __dmul_rn(vfac, vfac)); // search.cxx:464
    for (uint32_t g = 1 + threadIdx.x; g <= nchunks; g += blockDim.x) vs2[g] = acc;
}

// ScalePhase (search.cxx:565-568) on the group-major columns; optionally the un-scaled values (search.cxx:573-575,
// SURVEY Q3) go back to the original-order device arrays.
__global__ void k_scale_phase(double *__restrict__ x, double *__restrict__ y, double *__restrict__ z, double *__restrict__ vx,
                              double *__restrict__ vy, double *__restrict__ vz, const uint32_t *__restrict__ gid,
                              const uint32_t *__restrict__ idx, const double *__restrict__ vs2, uint32_t ntot, double xs,
                              double xsinv, double *__restrict__ pos_out, double *__restrict__ vel_out)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot) return;
    const double vsc = __ddiv_rn(1.0, sqrt(vs2[gid[s]]));
    const double vinv = __ddiv_rn(1.0, vsc);
    const double a = __dmul_rn(x[s], xs), b = __dmul_rn(y[s], xs), c = __dmul_rn(z[s], xs);
    const double d = __dmul_rn(vx[s], vsc), e = __dmul_rn(vy[s], vsc), f = __dmul_rn(vz[s], vsc);
    x[s] = a; y[s] = b; z[s] = c; vx[s] = d; vy[s] = e; vz[s] = f;
    if (pos_out) {
        const size_t i = idx[s];
        pos_out[3 * i] = __dmul_rn(a, xsinv); pos_out[3 * i + 1] = __dmul_rn(b, xsinv); pos_out[3 * i + 2] = __dmul_rn(c, xsinv);
        vel_out[3 * i] = __dmul_rn(d, vinv); vel_out[3 * i + 1] = __dmul_rn(e, vinv); vel_out[3 * i + 2] = __dmul_rn(f, vinv);
    }
}

// ---------------------------------------------------------------------------------------------------
// cells in scaled position space
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long dbl_ord(double d)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(d);
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}
static double ord_dbl_host(unsigned long long u)
{
    u = (u & 0x8000000000000000ull) ? (u & 0x7fffffffffffffffull) : ~u;
    double d;
    std::memcpy(&d, &u, 8);
    return d;
}

__global__ void k_minmax3(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z, uint32_t n,
                          unsigned long long *__restrict__ out /* min[3], max[3] as ordered ints */)
{
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double v[3] = {x[i], y[i], z[i]};
        for (int k = 0; k < 3; k++) { mn[k] = fmin(mn[k], v[k]); mx[k] = fmax(mx[k], v[k]); }
    }
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o; o >>= 1) {
            mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
        if ((threadIdx.x & 31) == 0) { atomicMin(&out[k], dbl_ord(mn[k])); atomicMax(&out[3 + k], dbl_ord(mx[k])); }
    }
}

struct CellOrigin { long long c0[3]; };

__global__ void k_keys6(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                        const uint32_t *__restrict__ gid, uint32_t ntot, CellOrigin o, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot) return;
    const uint64_t cx = (uint64_t)((long long)floor(x[s]) - o.c0[0]);
    const uint64_t cy = (uint64_t)((long long)floor(y[s]) - o.c0[1]);
    const uint64_t cz = (uint64_t)((long long)floor(z[s]) - o.c0[2]);
    keys[s] = ((uint64_t)gid[s] << (3 * CBITS)) | (cz << (2 * CBITS)) | (cy << CBITS) | cx;
    vals[s] = s;
}

__global__ void k_gather6(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                          const double *__restrict__ vx, const double *__restrict__ vy, const double *__restrict__ vz,
                          const uint32_t *__restrict__ ord, uint32_t ntot, double *__restrict__ q /* 6 columns of ntot */)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= ntot) return;
    const uint32_t s = ord[p];
    q[p] = x[s]; q[(size_t)ntot + p] = y[s]; q[2 * (size_t)ntot + p] = z[s];
    q[3 * (size_t)ntot + p] = vx[s]; q[4 * (size_t)ntot + p] = vy[s]; q[5 * (size_t)ntot + p] = vz[s];
}

__global__ void k_flag_heads6(const uint64_t *__restrict__ keys, uint32_t n, uint32_t *__restrict__ flags)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    flags[k] = (k == 0 || keys[k] != keys[k - 1]) ? 1u : 0u;
}

__global__ void k_fill_cells6(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl,
                              uint32_t n, uint32_t *__restrict__ cell_of, uint32_t *__restrict__ cell_start, uint64_t *__restrict__ cell_key,
                              uint32_t ncells)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t f = flags[k];
    const uint32_t c = excl[k] - (f ? 0u : 1u);
    cell_of[k] = c;
    if (f) { cell_start[c] = k; cell_key[c] = keys[k]; }
    if (k == n - 1) cell_start[ncells] = n;
}

__global__ void k_chunk_counts(const uint32_t *__restrict__ cell_start, uint32_t ncells, uint32_t *__restrict__ nchunk)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < ncells) nchunk[c] = (cell_start[c + 1] - cell_start[c] + 31) >> 5;
}

// chunk -> its cell (most cells are one chunk; a dense cell writes its few chunk entries in a loop)
__global__ void k_chunk_cells(const uint32_t *__restrict__ chunk_start, const uint32_t *__restrict__ nchunk, uint32_t ncells, uint32_t *__restrict__ chunk_cell)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    const uint32_t k0 = chunk_start[c], k1 = k0 + nchunk[c];
    for (uint32_t k = k0; k < k1; k++) chunk_cell[k] = c;
}

// ---------------------------------------------------------------------------------------------------
// linking
// ---------------------------------------------------------------------------------------------------
struct Link6 {
    const double *q;            // 6 columns of ntot (sorted order)
    uint32_t ntot;
    const uint32_t *cell_of, *cell_start;
    const uint64_t *cell_key;
    uint32_t ncells;
    const uint64_t *hkeys;
    const uint32_t *hvals;
    uint32_t hmask;
    uint32_t *parent;
    const uint32_t *uni;        // per cell: common root of all its particles, or NONE32
    const uint32_t *chunk_start; // exclusive scan of per-cell 32-particle chunk counts
    const uint32_t *chunk_cell;  // chunk -> cell
    uint32_t nchunks;
    unsigned long long *st;      // VRGPU_LINK6_STATS=1: counters (a measuring aid)
    unsigned int *next_item;     // work counter: chunks are handed out in order, so that the warps in flight work on neighbouring cells
    uint32_t sparse_a, sparse_b; // a cell of at most sparse_a particles whose neighbour cells hold at most sparse_b each takes the sparse path
    double p6, p7, ip6, ip7;    // criterion predicate only (FOF6d: params[6], params[7] and their reciprocals)
};

__device__ __forceinline__ bool links6(double ax, double ay, double az, double avx, double avy, double avz, double bx, double by,
                                       double bz, double bvx, double bvy, double bvz)
{
    double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by), dz = __dsub_rn(az, bz);
    double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
    dx = __dsub_rn(avx, bvx); dy = __dsub_rn(avy, bvy); dz = __dsub_rn(avz, bvz);
    double v2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
    return __dadd_rn(d2, v2) < 1.0;
}

// FOF6d (NBodylib/src/KDTree/FOFFunc.h:48-55), the comparison function of KDTree::FOFCriterion in the 6D core search
// (search.cxx:1607-1611): total += dx_j^2/params[6]; total += dv_j^2/params[7], j = 0..2 in this
// order, link iff total < 1.  The six IEEE divisions are only spent on pairs within 1e-9 of the threshold; everything
// else is decided by the same sum taken with reciprocals (which differs from the exact one by a few 1e-16).
__device__ __forceinline__ bool links6_fof6d(const Link6 &L, double ax, double ay, double az, double avx, double avy, double avz, double bx,
                                             double by, double bz, double bvx, double bvy, double bvz)
{
    const double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by), dz = __dsub_rn(az, bz);
    const double ux = __dsub_rn(avx, bvx), uy = __dsub_rn(avy, bvy), uz = __dsub_rn(avz, bvz);
    const double x2 = __dmul_rn(dx, dx), y2 = __dmul_rn(dy, dy), z2 = __dmul_rn(dz, dz);
    const double a2 = __dmul_rn(ux, ux), b2 = __dmul_rn(uy, uy), c2 = __dmul_rn(uz, uz);
    const double approx = (x2 + y2 + z2) * L.ip6 + (a2 + b2 + c2) * L.ip7;
    if (approx < 1.0 - 1e-9) return true;
    if (approx > 1.0 + 1e-9) return false;
    double total = 0.0;
    total = __dadd_rn(total, __ddiv_rn(x2, L.p6)); total = __dadd_rn(total, __ddiv_rn(a2, L.p7));
    total = __dadd_rn(total, __ddiv_rn(y2, L.p6)); total = __dadd_rn(total, __ddiv_rn(b2, L.p7));
    total = __dadd_rn(total, __ddiv_rn(z2, L.p6)); total = __dadd_rn(total, __ddiv_rn(c2, L.p7));
    return total < 1.0;
}
// CRIT = false: the scaled-coordinate predicate of the 6DFOF stage; CRIT = true: FOF6d on unscaled coordinates
template <bool CRIT>
__device__ __forceinline__ bool links6p(const Link6 &L, double ax, double ay, double az, double avx, double avy, double avz, double bx, double by,
                                        double bz, double bvx, double bvy, double bvz)
{
    if (CRIT) return links6_fof6d(L, ax, ay, az, avx, avy, avz, bx, by, bz, bvx, bvy, bvz);
    return links6(ax, ay, az, avx, avy, avz, bx, by, bz, bvx, bvy, bvz);
}

// seeding: every particle against (up to) the first 32 particles of its own cell that precede it; stop at the first link
template <bool CRIT = false> __global__ void __launch_bounds__(256) k_seed6(Link6 L)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L.ntot) return;
    const uint32_t a0 = L.cell_start[L.cell_of[i]];
    const uint32_t jend = min(i, a0 + 32u);
    if (jend <= a0) return;
    const size_t n = L.ntot;
    const double ax = L.q[i], ay = L.q[n + i], az = L.q[2 * n + i], avx = L.q[3 * n + i], avy = L.q[4 * n + i], avz = L.q[5 * n + i];
    for (uint32_t j = a0; j < jend; j++) {
        if (links6p<CRIT>(L, ax, ay, az, avx, avy, avz, L.q[j], L.q[n + j], L.q[2 * n + j], L.q[3 * n + j], L.q[4 * n + j], L.q[5 * n + j])) {
            uf_union(L.parent, i, j);
            return;
        }
    }
}

// one warp per cell (grid stride): common root of the cell or NONE32.  parent[] must be flat.
__global__ void __launch_bounds__(256) k_cell_uniform(const uint32_t *__restrict__ parent, const uint32_t *__restrict__ cell_start,
                                                      uint32_t ncells, uint32_t *__restrict__ uni)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t c = wg; c < ncells; c += nw) {
        const uint32_t a0 = cell_start[c], a1 = cell_start[c + 1];
        const uint32_t r0 = parent[a0];
        bool same = true;
        for (uint32_t p = a0 + lane; p < a1 && same; p += 32) same = parent[p] == r0;
        same = __all_sync(0xffffffffu, same);
        if (lane == 0) uni[c] = same ? r0 : NONE32;
    }
}

constexpr int L6_WARPS = 8;
constexpr uint32_t L6_SPARSE_A = 8, L6_SPARSE_B = 32; // a cell of at most 8 particles whose 14 neighbour cells hold at most 32 each

// exhaustive pass: one warp per 32-particle chunk of a cell; 13 forward neighbour cells + the cell itself.
template <bool CRIT = false> __global__ void __launch_bounds__(L6_WARPS * 32) k_link6(Link6 L)
{
    __shared__ double tq[L6_WARPS][6][32];
    __shared__ uint32_t tr[L6_WARPS][32];
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const size_t n = L.ntot;
    // Chunks are taken from a counter, not by a fixed stride: the chunks in flight at any moment are then a contiguous window of
    // cells (neighbours along x, the same rows of the same group), whose neighbour tiles overlap and stay in L2.  With a grid
    // stride over eight waves of blocks every wave swept the whole array: 33 GB of DRAM traffic at 512^3 for 2.6 GB of columns.
    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(L.next_item, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= L.nchunks) break;
        const uint32_t A = L.chunk_cell[item];
        const uint32_t a0 = L.cell_start[A], a1 = L.cell_start[A + 1];
        const uint32_t i = a0 + ((item - L.chunk_start[A]) << 5) + lane;
        const bool valid = i < a1;
        double ax = 0, ay = 0, az = 0, avx = 0, avy = 0, avz = 0;
        uint32_t ri = NONE32;
        if (valid) {
            ax = L.q[i]; ay = L.q[n + i]; az = L.q[2 * n + i]; avx = L.q[3 * n + i]; avy = L.q[4 * n + i]; avz = L.q[5 * n + i];
            ri = uf_find(L.parent, i);
        }
        const uint64_t key = L.cell_key[A];
        const uint32_t uniA = L.uni[A];
        // neighbour lookup: lane o < 14 owns half-space offset o (0 = the cell itself)
        int nb = -1;
        if (lane < 14) {
            int dx = 0, dy = 0, dz = 0;
            if (lane > 0) {
                // forward half space: dz=1 (9), dz=0 & dy=1 (3), dz=0 & dy=0 & dx=1 (1)
                int o = (int)lane - 1;
                if (o < 9) { dz = 1; dy = o / 3 - 1; dx = o % 3 - 1; }
                else if (o < 12) { dz = 0; dy = 1; dx = o - 9 - 1; }
                else { dz = 0; dy = 0; dx = 1; }
            }
            const long long cx = (long long)(key & CMASK) + dx, cy = (long long)((key >> CBITS) & CMASK) + dy,
                            cz = (long long)((key >> (2 * CBITS)) & CMASK) + dz;
            if (lane == 0) nb = (int)A;
            else if (cx >= 0 && cy >= 0 && cz >= 0 && cx <= (long long)CMASK && cy <= (long long)CMASK && cz <= (long long)CMASK) {
                const uint64_t nkey = (key & ~((1ull << (3 * CBITS)) - 1)) | ((uint64_t)cz << (2 * CBITS)) | ((uint64_t)cy << CBITS) | (uint64_t)cx;
                nb = hash_lookup(L.hkeys, L.hvals, L.hmask, nkey);
            }
        }
        unsigned found = __ballot_sync(0xffffffffu, nb >= 0);
        // Sparse neighbourhoods (the outskirts of every group: a handful of particles per cell) leave most lanes of a chunk idle
        // and pay one tile round trip per neighbour cell.  There the (particle, neighbour cell) pairs are spread over the lanes
        // instead, and every lane walks its neighbour cell straight from global memory.
        if (a1 - a0 <= L.sparse_a) {
            uint32_t nb0 = 0, nbsz = 0, nbuni = NONE32;
            if (nb >= 0) { nb0 = L.cell_start[nb]; nbsz = L.cell_start[nb + 1] - nb0; nbuni = L.uni[nb]; }
            if (!__any_sync(0xffffffffu, nbsz > L.sparse_b)) {
                const uint32_t ka = a1 - a0;
                const unsigned nitems = ka * (unsigned)__popc(found);
                for (unsigned base = 0; base < nitems; base += 32) {
                    const unsigned it = base + lane;
                    const bool on = it < nitems;
                    const unsigned il = on ? it % ka : 0u, oi = on ? it / ka : 0u;
                    const int src = (int)__fns(found, 0, (int)oi + 1);
                    const uint32_t B = (uint32_t)__shfl_sync(0xffffffffu, nb, src), b0 = __shfl_sync(0xffffffffu, nb0, src),
                                   bs = __shfl_sync(0xffffffffu, nbsz, src), uniB = __shfl_sync(0xffffffffu, nbuni, src);
                    if (!on || (uniA != NONE32 && uniA == uniB)) continue;
                    if (L.st) { atomicAdd(L.st + 4, 1ull); atomicAdd(L.st + 5, (unsigned long long)bs); }
                    const uint32_t pi = a0 + il;
                    const bool self = B == A;
                    const double px = L.q[pi], py = L.q[n + pi], pz = L.q[2 * n + pi], pvx = L.q[3 * n + pi], pvy = L.q[4 * n + pi], pvz = L.q[5 * n + pi];
                    uint32_t rp = uf_find(L.parent, pi);
                    for (uint32_t j = self ? pi + 1 : b0; j < b0 + bs; j++) {
                        if (((volatile uint32_t *)L.parent)[j] == rp) continue; // possibly stale ancestor: only costs a redundant test
                        if (links6p<CRIT>(L, px, py, pz, pvx, pvy, pvz, L.q[j], L.q[n + j], L.q[2 * n + j], L.q[3 * n + j], L.q[4 * n + j], L.q[5 * n + j])) {
                            uf_union(L.parent, pi, j);
                            if (uniB != NONE32 && !self) break; // B was one component already: nothing left to find in it
                            rp = uf_find(L.parent, pi);
                        }
                    }
                }
                continue;
            }
        }
        if (L.st && lane == 0) { atomicAdd(L.st + 6, 1ull); atomicAdd(L.st + 7, (unsigned long long)(a1 - a0 < 32u ? a1 - a0 : 32u)); }
        while (found) {
            const int src = __ffs(found) - 1;
            found &= found - 1;
            const uint32_t B = (uint32_t)__shfl_sync(0xffffffffu, nb, src);
            const uint32_t uniB = L.uni[B];
            if (uniA != NONE32 && uniA == uniB) continue; // both cells already inside one component
            const uint32_t b0 = L.cell_start[B], b1 = L.cell_start[B + 1];
            const bool self = B == A;
            if (valid) ri = uf_find(L.parent, i); // links made by other warps since
            for (uint32_t tb = b0; tb < b1; tb += 32) {
                // lanes already in B's (uniform) component have nothing left to find in B
                if (uniB != NONE32 && __all_sync(0xffffffffu, !valid || ri == uniB)) break;
                if (self && tb + 31 <= a0 + ((item - L.chunk_start[A]) << 5)) continue; // tile entirely before this chunk (j > i only)
                const uint32_t j = tb + lane;
                // The dense cells at the centre of a group are one component long before this pass ends, yet a few stragglers keep
                // them from being `uniform`.  So first the tile's current roots alone: a tile whose 32 particles share one root that
                // every particle of the chunk has too needs neither its coordinates nor the 32-step loop.
                const uint32_t rj = j < b1 ? uf_find(L.parent, j) : NONE32;
                const uint32_t r0 = __shfl_sync(0xffffffffu, rj, 0);
                if (__all_sync(0xffffffffu, (j >= b1 || rj == r0) && (!valid || ri == r0))) {
                    if (L.st && lane == 0) atomicAdd(L.st + 1, 1ull);
                    continue;
                }
                __syncwarp();
                if (j < b1) {
                    tq[w][0][lane] = L.q[j]; tq[w][1][lane] = L.q[n + j]; tq[w][2][lane] = L.q[2 * n + j];
                    tq[w][3][lane] = L.q[3 * n + j]; tq[w][4][lane] = L.q[4 * n + j]; tq[w][5][lane] = L.q[5 * n + j];
                    // the CURRENT root (with path halving), not parent[j]: once components have met, a stale ancestor would send every
                    // pair of the two cells through the full test and a union that finds nothing to do
                    tr[w][lane] = rj;
                }
                __syncwarp();
                const int cnt = (int)min(32u, b1 - tb);
                if (L.st && lane == 0) atomicAdd(L.st + 0, 1ull);
                if (L.st) atomicAdd(L.st + 2, valid ? (unsigned long long)cnt : 0ull);
                if (valid) {
                    for (int jj = 0; jj < cnt; jj++) {
                        const uint32_t jx = tb + jj;
                        if (self && jx <= i) continue;
                        if (tr[w][jj] == ri) continue;
                        if (L.st) atomicAdd(L.st + 3, 1ull);
                        if (links6p<CRIT>(L, ax, ay, az, avx, avy, avz, tq[w][0][jj], tq[w][1][jj], tq[w][2][jj], tq[w][3][jj], tq[w][4][jj], tq[w][5][jj])) {
                            uf_union(L.parent, i, jx);
                            ri = uf_find(L.parent, i);
                        }
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// labels
// ---------------------------------------------------------------------------------------------------
// particles are in (group, cell) order, so a warp mostly holds members of one component: one atomic per (warp, root)
__global__ void k_comp_stats(const uint32_t *__restrict__ root, const uint32_t *__restrict__ ord, const uint32_t *__restrict__ idx,
                             const long long *__restrict__ pid, uint32_t ntot, uint32_t *__restrict__ csize,
                             unsigned long long *__restrict__ cminpid)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = p < ntot;
    uint32_t r = 0xffffffffu;
    unsigned long long v = ~0ull;
    if (valid) {
        r = root[p];
        const uint32_t i = idx[ord[p]];
        v = pid ? (unsigned long long)pid[i] : (unsigned long long)i;
    }
    const unsigned peers = __match_any_sync(0xffffffffu, r);
    const unsigned long long mn = warp_min_u64(peers, v);
    if (valid && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) {
        atomicAdd(&csize[r], (uint32_t)__popc(peers));
        atomicMin(&cminpid[r], mn);
    }
}

__global__ void k_write_pfof6(const uint32_t *__restrict__ root, const uint32_t *__restrict__ ord, const uint32_t *__restrict__ idx,
                              const uint32_t *__restrict__ gid_of, uint32_t ntot, long long *__restrict__ pfof)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= ntot) return;
    pfof[idx[ord[p]]] = (long long)gid_of[root[p]];
}

__global__ void k_remap_chunks(const long long *__restrict__ pfof, size_t n, long long iend, long long *__restrict__ out)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long g = pfof[i];
    out[i] = g > iend ? iend : g;
}

__global__ void k_first_ungrouped(const long long *__restrict__ pfof, size_t n, unsigned int *__restrict__ first)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && pfof[i] == 0) atomicMin(first, (unsigned int)i);
}

// binning coordinates of the criterion search
__global__ void k_bin_coords(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z, uint32_t n, double inv,
                             double *__restrict__ bx, double *__restrict__ by, double *__restrict__ bz)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    bx[s] = x[s] * inv; by[s] = y[s] * inv; bz[s] = z[s] * inv;
}

// Back half shared by the 6DFOF stage and the criterion search: cells of edge 1 in the BINNING coordinates, particle-level
// union-find (seed pass, uniform-cell flags, exhaustive pass), canonical labels.  crit == nullptr: gs holds the scaled
// phase-space columns, which are both the binning coordinates and the predicate's input.  crit != nullptr ({params[6],
// params[7]}): gs holds unscaled columns, the binning coordinates are x / (sqrt(params[6]) (1 + 1e-9)) -- a linked pair has
// |dx_j| < sqrt(params[6]) on every axis, so it lies in adjacent cells -- and the predicate is FOF6d.
static long long link6_back(vrgpu_ctx *ctx, const DeviceParticles &P, GroupSet &gs, uint32_t G, long long minsize, long long *pfof6d_dev,
                            const double *crit)
{
    const int T = 256;
    const uint32_t ntot = (uint32_t)gs.ntot;
    DBuf<double> bx, by, bz;
    const double *kx = gs.x.p, *ky = gs.y.p, *kz = gs.z.p;
    if (crit) {
        bx.alloc(ctx, ntot); by.alloc(ctx, ntot); bz.alloc(ctx, ntot);
        const double inv = 1.0 / (std::sqrt(crit[0]) * (1.0 + 1e-9));
        VR_LAUNCH(ctx, k_bin_coords, grid_for(ntot, T), T, 0, gs.x.p, gs.y.p, gs.z.p, ntot, inv, bx.p, by.p, bz.p);
        kx = bx.p; ky = by.p; kz = bz.p;
    }
    // ---- cells of edge 1 in scaled position space ----
    DBuf<uint64_t> keys(ctx, ntot), keys_tmp(ctx, ntot);
    DBuf<uint32_t> ord(ctx, ntot), ord_tmp(ctx, ntot);
    DBuf<double> q(ctx, 6 * (size_t)ntot);
    DBuf<uint32_t> cell_of(ctx, ntot), cell_start, nchunk, chunk_start, chunk_cell;
    DBuf<uint64_t> cell_key;
    uint32_t ncells = 0, nwork = 0;
    {
        KScope k(ctx, "fof6d.cells");
        DBuf<unsigned long long> mm(ctx, 6);
        unsigned long long init[6] = {~0ull, ~0ull, ~0ull, 0ull, 0ull, 0ull};
        VR_CUDA(cudaMemcpyAsync(mm.p, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
        VR_LAUNCH(ctx, k_minmax3, std::min<unsigned>(grid_for(ntot, T), (unsigned)ctx->sm_count * 8u), T, 0, kx, ky, kz, ntot, mm.p);
        unsigned long long enc[6];
        VR_CUDA(cudaMemcpyAsync(enc, mm.p, sizeof(enc), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        CellOrigin o{};
        for (int a = 0; a < 3; a++) {
            const double lo = ord_dbl_host(enc[a]), hi = ord_dbl_host(enc[3 + a]);
            VR_REQUIRE(std::isfinite(lo) && std::isfinite(hi), VRGPU_EINVAL, "non-finite scaled positions in 6DFOF (zero velocity dispersion?)");
            o.c0[a] = (long long)std::floor(lo);
            VR_REQUIRE((long long)std::floor(hi) - o.c0[a] <= (long long)CMASK, VRGPU_EUNSUPPORTED,
                       "6DFOF: extent / linking length exceeds 16384 cells per axis");
        }
        VR_REQUIRE(bits_for_u64v(G) + 3 * CBITS <= 64, VRGPU_EUNSUPPORTED, "6DFOF: too many 3D groups for the cell key");
        VR_LAUNCH(ctx, k_keys6, grid_for(ntot, T), T, 0, kx, ky, kz, gs.gid.p, ntot, o, keys.p, ord.p);
        radix_sort_pairs(ctx, keys.p, ord.p, keys_tmp.p, ord_tmp.p, ntot, bits_for_u64v(G) + 3 * CBITS);
        VR_LAUNCH(ctx, k_gather6, grid_for(ntot, T), T, 0, gs.x.p, gs.y.p, gs.z.p, gs.vx.p, gs.vy.p, gs.vz.p, ord.p, ntot, q.p);
        DBuf<uint32_t> flags(ctx, ntot), excl(ctx, ntot), total(ctx, 1);
        VR_LAUNCH(ctx, k_flag_heads6, grid_for(ntot, T), T, 0, keys.p, ntot, flags.p);
        exclusive_scan_u32(ctx, flags.p, excl.p, ntot, total.p);
        VR_CUDA(cudaMemcpyAsync(&ncells, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        cell_start.alloc(ctx, (size_t)ncells + 1);
        cell_key.alloc(ctx, ncells);
        VR_LAUNCH(ctx, k_fill_cells6, grid_for(ntot, T), T, 0, keys.p, flags.p, excl.p, ntot, cell_of.p, cell_start.p, cell_key.p, ncells);
        nchunk.alloc(ctx, ncells);
        chunk_start.alloc(ctx, ncells);
        VR_LAUNCH(ctx, k_chunk_counts, grid_for(ncells, T), T, 0, cell_start.p, ncells, nchunk.p);
        exclusive_scan_u32(ctx, nchunk.p, chunk_start.p, ncells, total.p);
        VR_CUDA(cudaMemcpyAsync(&nwork, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        chunk_cell.alloc(ctx, (size_t)nwork + 1);
        VR_LAUNCH(ctx, k_chunk_cells, grid_for(ncells, T), T, 0, chunk_start.p, nchunk.p, ncells, chunk_cell.p);
    }
    keys.reset(); keys_tmp.reset(); ord_tmp.reset();
    bx.reset(); by.reset(); bz.reset();
    // the group-major columns are no longer needed (q holds the sorted copy)
    gs.x.reset(); gs.y.reset(); gs.z.reset(); gs.vx.reset(); gs.vy.reset(); gs.vz.reset(); gs.m.reset();

    // ---- union-find over particles ----
    DBuf<uint32_t> parent(ctx, ntot), root(ctx, ntot), uni(ctx, ncells);
    {
        size_t h = 1;
        while (h < 2 * (size_t)ncells) h <<= 1;
        DBuf<uint64_t> hkeys(ctx, h);
        DBuf<uint32_t> hvals(ctx, h);
        Link6 L{};
        L.q = q.p; L.ntot = ntot; L.cell_of = cell_of.p; L.cell_start = cell_start.p; L.cell_key = cell_key.p; L.ncells = ncells;
        L.hkeys = hkeys.p; L.hvals = hvals.p; L.hmask = (uint32_t)(h - 1); L.parent = parent.p; L.uni = uni.p;
        L.chunk_start = chunk_start.p; L.chunk_cell = chunk_cell.p; L.nchunks = nwork;
        static const int sparse_a = [] { const char *e = getenv("VRGPU_LINK6_SPARSE"); return e ? atoi(e) : (int)L6_SPARSE_A; }();
        L.sparse_a = (uint32_t)sparse_a;
        DBuf<unsigned int> next_item(ctx, 1);
        VR_CUDA(cudaMemsetAsync(next_item.p, 0, sizeof(unsigned int), ctx->stream));
        L.next_item = next_item.p;
        static const bool link6_stats = [] { const char *e = getenv("VRGPU_LINK6_STATS"); return e && atoi(e) == 1; }();
        DBuf<unsigned long long> lst;
        if (link6_stats) { lst.alloc(ctx, 8); VR_CUDA(cudaMemsetAsync(lst.p, 0, 64, ctx->stream)); L.st = lst.p; }
        static const int sparse_b = [] { const char *e = getenv("VRGPU_LINK6_SPARSE_B"); return e ? atoi(e) : (int)L6_SPARSE_B; }();
        L.sparse_b = (uint32_t)sparse_b;
        if (crit) { L.p6 = crit[0]; L.p7 = crit[1]; L.ip6 = 1.0 / crit[0]; L.ip7 = 1.0 / crit[1]; }
        {
            KScope k(ctx, "fof6d.hash_build");
            VR_LAUNCH(ctx, k_hash_fill, grid_for(h, T), T, 0, hkeys.p, h);
            VR_LAUNCH(ctx, k_hash_insert, grid_for(ncells, T), T, 0, cell_key.p, ncells, hkeys.p, hvals.p, (uint32_t)(h - 1), (uint32_t *)nullptr);
            VR_LAUNCH(ctx, k_uf_init, grid_for(ntot, T), T, 0, parent.p, ntot);
        }
        {
            KScope k(ctx, "fof6d.seed");
            if (crit) VR_LAUNCH(ctx, k_seed6<true>, grid_for(ntot, T), T, 0, L);
            else VR_LAUNCH(ctx, k_seed6<false>, grid_for(ntot, T), T, 0, L);
            VR_LAUNCH(ctx, k_uf_flatten, grid_for(ntot, T), T, 0, parent.p, ntot, parent.p);
            const unsigned grid = std::max(1u, std::min<unsigned>(grid_for((size_t)ncells * 32, T), (unsigned)ctx->sm_count * 16u));
            VR_LAUNCH(ctx, k_cell_uniform, grid, T, 0, parent.p, cell_start.p, ncells, uni.p);
        }
        {
            KScope k(ctx, "fof6d.link");
            const unsigned grid = std::max(1u, std::min<unsigned>((nwork + L6_WARPS - 1) / L6_WARPS, (unsigned)ctx->sm_count * 32u));
            if (crit) VR_LAUNCH(ctx, k_link6<true>, grid, L6_WARPS * 32, 0, L);
            else VR_LAUNCH(ctx, k_link6<false>, grid, L6_WARPS * 32, 0, L);
            if (link6_stats) {
                unsigned long long h[8];
                VR_CUDA(cudaMemcpyAsync(h, lst.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
                VR_CUDA(cudaStreamSynchronize(ctx->stream));
                fprintf(stderr, "[link6 stats] particles %u cells %u chunks %u | dense chunks %llu (particles %llu), tiles %llu (+ %llu skipped on their roots), valid-lane iterations %llu, "
                        "full tests %llu | sparse items %llu, their neighbour particles %llu\n", ntot, ncells, nwork, h[6], h[7], h[0], h[1], h[2], h[3], h[4], h[5]);
            }
            VR_LAUNCH(ctx, k_uf_flatten, grid_for(ntot, T), T, 0, parent.p, ntot, root.p);
        }
    }

    // ---- labels: (size desc, min pid asc) over all chunks ----
    uint32_t ng = 0;
    {
        KScope k(ctx, "fof6d.labels");
        DBuf<uint32_t> csize(ctx, ntot), gid_of(ctx, ntot);
        DBuf<unsigned long long> cminpid(ctx, ntot);
        VR_CUDA(cudaMemsetAsync(csize.p, 0, sizeof(uint32_t) * ntot, ctx->stream));
        VR_CUDA(cudaMemsetAsync(cminpid.p, 0xff, sizeof(unsigned long long) * ntot, ctx->stream));
        VR_LAUNCH(ctx, k_comp_stats, grid_for(ntot, T), T, 0, root.p, ord.p, gs.idx.p, P.has_pid ? P.pid.p : (const long long *)nullptr, ntot,
                  csize.p, cminpid.p);
        uint32_t ms = minsize < 1 ? 1u : (minsize > 0xffffffffLL ? 0xffffffffu : (uint32_t)minsize);
        ng = order_groups(ctx, root.p, csize.p, cminpid.p, ntot, ms, gid_of.p);
        VR_LAUNCH(ctx, k_write_pfof6, grid_for(ntot, T), T, 0, root.p, ord.p, gs.idx.p, gid_of.p, ntot, pfof6d_dev);
    }
    return (long long)ng;
}

long long fof6d_device(vrgpu_ctx *ctx, DeviceParticles &P, const long long *pfof3d_dev, long long ngroups3d, double b6, double vfac,
                       int mode, long long minsize, long long *pfof6d_dev, bool apply_q3)
{
    const int T = 256;
    VR_REQUIRE(P.has_vel, VRGPU_EINVAL, "6DFOF needs velocities");
    VR_REQUIRE(b6 > 0 && vfac > 0, VRGPU_EINVAL, "6DFOF linking lengths must be positive");
    VR_CUDA(cudaMemsetAsync(pfof6d_dev, 0, sizeof(long long) * (P.n ? P.n : 1), ctx->stream));
    if (P.n == 0 || ngroups3d <= 0) return 0;

    // ---- search chunks: one per 3D group; FOF6D (mode 4) merges every group with n <= ompunbindnum(1000) into ONE
    //      chunk (search.cxx:529-535) ----
    GroupSet gs;
    DBuf<long long> chunk_of;
    long long nchunks = ngroups3d;
    long long n_first = 0, inext = -1;
    {
        KScope k(ctx, "fof6d.regroup");
        if (mode == 4) {
            GroupSet g0;
            build_groupset(ctx, pfof3d_dev, P.n, ngroups3d, g0);
            std::vector<uint32_t> off((size_t)ngroups3d + 2);
            VR_CUDA(cudaMemcpyAsync(off.data(), g0.off.p, off.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
            long long iend = ngroups3d;
            for (long long g = 1; g <= ngroups3d; g++)
                if (off[g + 1] - off[g] <= 1000u) { iend = g; break; }
            n_first = off[2] - off[1];
            if (ngroups3d >= 2) {
                uint32_t h = 0;
                VR_CUDA(cudaMemcpyAsync(&h, g0.idx.p + off[2], sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
                VR_CUDA(cudaStreamSynchronize(ctx->stream));
                inext = h;
            } else {
                DBuf<unsigned int> first(ctx, 1);
                VR_CUDA(cudaMemsetAsync(first.p, 0xff, sizeof(unsigned int), ctx->stream));
                VR_LAUNCH(ctx, k_first_ungrouped, grid_for(P.n, T), T, 0, pfof3d_dev, P.n, first.p);
                unsigned int h = 0;
                VR_CUDA(cudaMemcpyAsync(&h, first.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
                VR_CUDA(cudaStreamSynchronize(ctx->stream));
                inext = h == 0xffffffffu ? 0 : (long long)h;
            }
            nchunks = iend;
            chunk_of.alloc(ctx, P.n);
            VR_LAUNCH(ctx, k_remap_chunks, grid_for(P.n, T), T, 0, pfof3d_dev, P.n, iend, chunk_of.p);
            build_groupset(ctx, chunk_of.p, P.n, nchunks, gs);
        } else {
            build_groupset(ctx, pfof3d_dev, P.n, ngroups3d, gs);
        }
    }
    if (gs.ntot == 0) return 0;
    { KScope k(ctx, "fof6d.gather"); gather_groupset(ctx, P, gs, true, true); }
    const uint32_t ntot = (uint32_t)gs.ntot;
    const uint32_t G = (uint32_t)nchunks;

    // ---- velocity scale per chunk ----
    DBuf<double> vs2(ctx, (size_t)G + 2);
    {
        KScope k(ctx, "fof6d.dispersion");
        // groups of at least this many members may go through the many-window path (measured on the 2e7-particle halo:
        // 292 -> 107 ms, profiles/r02a_probe.md); VRGPU_DISP_MULTI_MIN=0 switches it off
        long long multi_min = 65536;
        if (const char *e = getenv("VRGPU_DISP_MULTI_MIN")) multi_min = atoll(e);
        SeqScratch sc;
        DBuf<double> seqr;
        auto prepare_multi = [&](unsigned long long nmax) {
            sc.cap = (unsigned)((3ull * nmax + SQ_W - 1) / SQ_W);
            sc.wsum.alloc(ctx, sc.cap); sc.start.alloc(ctx, sc.cap); sc.spec.alloc(ctx, sc.cap); sc.nslow.alloc(ctx, 1);
            seqr.alloc(ctx, 8);
            VR_CUDA(cudaMemsetAsync(sc.nslow.p, 0, sizeof(unsigned int), ctx->stream));
            VR_CUDA(cudaMemsetAsync(seqr.p, 0, 8 * sizeof(double), ctx->stream));
        };
        if (mode == 4) {
            double mnorm = 1.0;
            if (P.has_mass) {
                VR_CUDA(cudaMemcpyAsync(&mnorm, P.mass.p + (inext >= 0 ? inext : 0), sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
                VR_CUDA(cudaStreamSynchronize(ctx->stream));
            }
            // group 1 is chunk 1 unless it was itself merged (then its members are no longer contiguous: use the 3D layout)
            if ((nchunks >= 2 || ngroups3d == 1) && multi_min > 0 && (long long)n_first >= multi_min) {
                prepare_multi((unsigned long long)n_first);
                dispersion_multi(ctx, sc, seqr, gs.vx.p, gs.vy.p, gs.vz.p, gs.m.p, (unsigned long long)n_first, vfac, true, mnorm, 1u, G, vs2.p);
            } else if (nchunks >= 2 || ngroups3d == 1) {
                VR_LAUNCH(ctx, k_single_dispersion, 1, SQ_NT, 0, gs.vx.p, gs.vy.p, gs.vz.p, gs.m.p, (uint32_t)n_first, mnorm, vfac, G, vs2.p);
            } else {
                GroupSet g0;
                build_groupset(ctx, pfof3d_dev, P.n, ngroups3d, g0);
                gather_groupset(ctx, P, g0, true, true);
                VR_LAUNCH(ctx, k_single_dispersion, 1, SQ_NT, 0, g0.vx.p, g0.vy.p, g0.vz.p, g0.m.p, (uint32_t)n_first, mnorm, vfac, G, vs2.p);
                VR_CUDA(cudaStreamSynchronize(ctx->stream));
            }
        } else {
            const unsigned grid = std::max(1u, std::min<unsigned>(grid_for((size_t)G * 32, T), (unsigned)ctx->sm_count * 8u));
            // the groups come in size order: the leading ones (at most 64) may take the many-window path
            uint32_t nskip = 0;
            bool big_done = false;
            static const int nside_streams = [] { const char *e = getenv("VRGPU_DISP_STREAMS"); return e ? std::max(0, std::min(4, atoi(e))) : 4; }();
            if (multi_min > 0) {
                const uint32_t look = std::min<uint32_t>(G, 64u);
                std::vector<uint32_t> hoff((size_t)look + 2);
                VR_CUDA(cudaMemcpyAsync(hoff.data(), gs.off.p, sizeof(uint32_t) * ((size_t)look + 2), cudaMemcpyDeviceToHost, ctx->stream));
                VR_CUDA(cudaStreamSynchronize(ctx->stream));
                // The one-CTA-per-group kernel runs all remaining groups at once, so it costs what its LARGEST group costs
                // (14.6 ns per member); the many-window path runs its groups one after the other (5.3 ns per member + ~0.15 ms
                // of launches each).  Take the prefix that minimises the sum.
                uint32_t elig = 0;
                while (elig < look && (long long)(hoff[elig + 2] - hoff[elig + 1]) >= multi_min) elig++;
                double best = 14.6e-6 * (double)(hoff[2] - hoff[1]), acc = 0.0;
                for (uint32_t k = 1; k <= elig; k++) {
                    // (on side streams the groups overlap: about half their serial cost is hidden -- profiles/r02_dispersion.md)
                    acc += (5.3e-6 * (double)(hoff[k + 1] - hoff[k]) + 0.15) * (nside_streams >= 2 ? 0.5 : 1.0);
                    const double rest = k < look ? 14.6e-6 * (double)(hoff[k + 2] - hoff[k + 1]) : 0.0;
                    if (acc + rest < best) { best = acc + rest; nskip = k; }
                }
                if (nskip) {
                    // The groups of the many-window path are independent of each other and of the one-CTA-per-group kernel: they go
                    // round robin to side streams (each with its own scratch), the kernel for the remaining groups runs on the
                    // context's stream meanwhile, and the context's stream waits for the side streams at the end.  The serial parts of
                    // a group (window prefix, the walker) then overlap with the parallel parts of the others.
                    const int NS = (int)std::min<uint32_t>((uint32_t)nside_streams, nskip);
                    if (NS >= 2) {
                        SeqScratch scs[4];
                        DBuf<double> rs[4];
                        cudaEvent_t ready = KScope::get(ctx), done[4] = {nullptr, nullptr, nullptr, nullptr};
                        VR_REQUIRE(ready != nullptr, VRGPU_ECUDA, "event creation failed");
                        for (int k = 0; k < NS; k++) {
                            if (!ctx->aux_stream[k]) VR_CUDA(cudaStreamCreateWithFlags(&ctx->aux_stream[k], cudaStreamNonBlocking));
                            const unsigned long long nmax = hoff[k + 2] - hoff[k + 1]; // the largest group this stream gets (sizes descend)
                            scs[k].cap = (unsigned)((3ull * nmax + SQ_W - 1) / SQ_W);
                            scs[k].wsum.alloc(ctx, scs[k].cap); scs[k].start.alloc(ctx, scs[k].cap); scs[k].spec.alloc(ctx, scs[k].cap);
                            scs[k].nslow.alloc(ctx, 1);
                            rs[k].alloc(ctx, 8);
                            VR_CUDA(cudaMemsetAsync(scs[k].nslow.p, 0, sizeof(unsigned int), ctx->stream));
                            VR_CUDA(cudaMemsetAsync(rs[k].p, 0, 8 * sizeof(double), ctx->stream));
                            done[k] = KScope::get(ctx);
                            VR_REQUIRE(done[k] != nullptr, VRGPU_ECUDA, "event creation failed");
                        }
                        VR_CUDA(cudaEventRecord(ready, ctx->stream)); // the columns and the cleared scratch
                        cudaStream_t main_stream = ctx->stream;
                        struct Restore { vrgpu_ctx *c; cudaStream_t s; ~Restore() { c->stream = s; } } restore{ctx, main_stream};
                        for (int k = 0; k < NS; k++) VR_CUDA(cudaStreamWaitEvent(ctx->aux_stream[k], ready, 0));
                        for (uint32_t g = 1; g <= nskip; g++) {
                            const int k = (int)((g - 1) % (uint32_t)NS);
                            const uint32_t s0 = hoff[g];
                            ctx->stream = ctx->aux_stream[k]; // the launches of dispersion_multi go to the side stream
                            dispersion_multi(ctx, scs[k], rs[k], gs.vx.p + s0, gs.vy.p + s0, gs.vz.p + s0, gs.m.p + s0, (unsigned long long)(hoff[g + 1] - s0),
                                             vfac, false, 0.0, g, 1u, vs2.p);
                            ctx->stream = main_stream;
                        }
                        for (int k = 0; k < NS; k++) VR_CUDA(cudaEventRecord(done[k], ctx->aux_stream[k]));
                        VR_LAUNCH(ctx, k_group_dispersion_big, G, SQ_NT, 0, gs.vx.p, gs.vy.p, gs.vz.p, gs.m.p, gs.off.p, vfac, vs2.p, nskip);
                        VR_LAUNCH(ctx, k_group_dispersion, grid, T, 0, gs.vx.p, gs.vy.p, gs.vz.p, gs.m.p, gs.off.p, G, vfac, vs2.p);
                        for (int k = 0; k < NS; k++) VR_CUDA(cudaStreamWaitEvent(ctx->stream, done[k], 0));
                        // the scratch goes back to the pool (ordered on the context's stream) only behind those waits
                        VR_CUDA(cudaStreamSynchronize(ctx->stream));
                        ctx->evpool.push_back(ready);
                        for (int k = 0; k < NS; k++) ctx->evpool.push_back(done[k]);
                        big_done = true;
                    } else {
                        prepare_multi((unsigned long long)(hoff[2] - hoff[1]));
                        for (uint32_t g = 1; g <= nskip; g++) {
                            const uint32_t s0 = hoff[g];
                            dispersion_multi(ctx, sc, seqr, gs.vx.p + s0, gs.vy.p + s0, gs.vz.p + s0, gs.m.p + s0, (unsigned long long)(hoff[g + 1] - s0),
                                             vfac, false, 0.0, g, 1u, vs2.p);
                        }
                    }
                }
            }
            if (!big_done) {
                VR_LAUNCH(ctx, k_group_dispersion_big, G, SQ_NT, 0, gs.vx.p, gs.vy.p, gs.vz.p, gs.m.p, gs.off.p, vfac, vs2.p, nskip);
                VR_LAUNCH(ctx, k_group_dispersion, grid, T, 0, gs.vx.p, gs.vy.p, gs.vz.p, gs.m.p, gs.off.p, G, vfac, vs2.p);
            }
        }
    }

    // ---- scale (and write the un-scaled values back, Q3) ----
    const double xs = 1.0 / b6;
    const double xsinv = 1.0 / xs;
    {
        KScope k(ctx, "fof6d.scale_phase");
        VR_LAUNCH(ctx, k_scale_phase, grid_for(ntot, T), T, 0, gs.x.p, gs.y.p, gs.z.p, gs.vx.p, gs.vy.p, gs.vz.p, gs.gid.p, gs.idx.p, vs2.p,
                  ntot, xs, xsinv, apply_q3 ? P.pos.p : (double *)nullptr, apply_q3 ? P.vel.p : (double *)nullptr);
    }

    return link6_back(ctx, P, gs, G, minsize, pfof6d_dev, nullptr);
}

// ---------------------------------------------------------------------------------------------------
// KDTree::FOFCriterion(FOF6d, params, numgroups, minsize, order = 1, ipcheckflag = 1, FOFcheckbg): the linking step of the
// 6D core search (search.cxx:1607-1611, 1664; KDFOF.cxx:133-240).  Particles whose check value
// is >= check_below are left out (FOFcheckbg, fofalgo.cxx:205-208); the others are the single "group" handed to the
// particle-level union-find with the FOF6d predicate.  The reference's tree only prunes with a lower bound of the
// position part of the sum, so its groups are the connected components of the predicate.
// ---------------------------------------------------------------------------------------------------
__global__ void k_check_mask(const double *__restrict__ check, size_t n, double below, long long *__restrict__ one)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) one[i] = (!check || check[i] < below) ? 1ll : 0ll;
}

long long fof_criterion6d_device(vrgpu_ctx *ctx, const DeviceParticles &P, const double *check_dev, double check_below, double p6, double p7,
                                 long long minsize, long long *pfof_dev)
{
    VR_REQUIRE(p6 > 0 && p7 > 0 && std::isfinite(p6) && std::isfinite(p7), VRGPU_EINVAL, "linking lengths must be positive and finite");
    VR_REQUIRE(P.has_vel, VRGPU_EINVAL, "the 6D criterion needs velocities");
    const int T = 256;
    VR_CUDA(cudaMemsetAsync(pfof_dev, 0, sizeof(long long) * (P.n ? P.n : 1), ctx->stream));
    if (P.n == 0) return 0;
    DBuf<long long> one(ctx, P.n);
    VR_LAUNCH(ctx, k_check_mask, grid_for(P.n, T), T, 0, check_dev, P.n, check_below, one.p);
    GroupSet gs;
    build_groupset(ctx, one.p, P.n, 1, gs);
    if (gs.ntot == 0) return 0;
    gather_groupset(ctx, P, gs, true, false);
    const double crit[2] = {p6, p7};
    return link6_back(ctx, P, gs, 1u, minsize, pfof_dev, crit);
}

// ---------------------------------------------------------------------------------------------------
// host write-back of the phase-space coordinates the path rewrote (AdjustStructureForPeriod + Q3)
// ---------------------------------------------------------------------------------------------------
__global__ void k_flag_pos(const long long *__restrict__ pfof, size_t n, uint32_t *__restrict__ flags)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = pfof[i] > 0 ? 1u : 0u;
}
__global__ void k_pack_phase(const double *__restrict__ pos, const double *__restrict__ vel, const uint32_t *__restrict__ flags,
                             const uint32_t *__restrict__ excl, size_t n, double *__restrict__ out /* 7 words per record */)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !flags[i]) return;
    const size_t s = excl[i];
    out[7 * s] = __longlong_as_double((long long)i);
    for (int k = 0; k < 3; k++) { out[7 * s + 1 + k] = pos[3 * i + k]; out[7 * s + 4 + k] = vel[3 * i + k]; }
}

// zero-copy variant: when the caller's records are page-locked (pinned / registered and mapped), the device stores the
// updated coordinates straight into them over PCIe -- no staging copy and no host scatter loop
__global__ void k_writeback_direct(unsigned char *__restrict__ base, size_t stride, size_t off_pos, size_t off_vel,
                                   const double *__restrict__ pos, const double *__restrict__ vel, const long long *__restrict__ pfof, size_t n)
{
    // six consecutive lanes store the six words of one record: when velocity follows position in the record (NBody::Particle)
    // the 48 bytes leave as one or two PCIe writes instead of six
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t i = t / 6;
    const unsigned w = (unsigned)(t - 6 * i);
    if (i >= n || pfof[i] <= 0) return;
    if (w < 3) *reinterpret_cast<double *>(base + i * stride + off_pos + 8 * w) = pos[3 * i + w];
    else if (vel) *reinterpret_cast<double *>(base + i * stride + off_vel + 8 * (w - 3)) = vel[3 * i + (w - 3)];
}

// device-visible alias of a page-locked host pointer, or nullptr for pageable memory
void *mapped_host_pointer(const void *p)
{
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (at.type == cudaMemoryTypeHost && at.devicePointer) return at.devicePointer;
    return nullptr;
}

// EXPERIMENT (VRGPU_ASYNC_WRITEBACK=1, run_search in api.cu): the zero-copy write-back of the 6DFOF coordinates on the copy
// stream, so that its PCIe stores (204 ms at 512^3, 26 ms at 256^3: profiles/r01f) run beside the unbinding stage instead
// of in front of it.  Returns the event that marks its end (nullptr when the caller's memory is not page-locked: the
// caller then takes the synchronous path); the inputs (P.pos, P.vel, pfof_dev) must stay untouched until that event.
cudaEvent_t writeback_phase_async(vrgpu_ctx *ctx, const vrgpu_view &view, const DeviceParticles &P, const long long *pfof_dev)
{
    if (P.n == 0) return nullptr;
    VR_REQUIRE(P.has_vel && view.off_vel != VRGPU_NOFIELD, VRGPU_EINVAL, "write-back needs velocities in the view");
    void *dev = mapped_host_pointer(view.base);
    if (!dev) return nullptr;
    if (!ctx->copy_stream) VR_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    cudaEvent_t ready = KScope::get(ctx), a = KScope::get(ctx), b = KScope::get(ctx);
    VR_CUDA(cudaEventRecord(ready, ctx->stream));              // everything that produced pos / vel / pfof
    VR_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ready, 0));
    VR_CUDA(cudaEventRecord(a, ctx->copy_stream));
    k_writeback_direct<<<grid_for(6 * P.n, 256), 256, 0, ctx->copy_stream>>>(static_cast<unsigned char *>(dev), view.stride, view.off_pos,
                                                                              view.off_vel, P.pos.p, P.vel.p, pfof_dev, P.n);
    ctx->stats.kernel_launches++;
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaEventRecord(b, ctx->copy_stream));
    cudaEvent_t done = KScope::get(ctx);
    VR_CUDA(cudaEventRecord(done, ctx->copy_stream));
    if (ctx->ktiming) ctx->kpending.push_back({"io.writeback", a, b}); // recycled by resolve_ktimes
    else { ctx->evpool.push_back(a); ctx->evpool.push_back(b); }
    ctx->evpool.push_back(ready);
    ctx->stats.d2h_bytes += (long long)ctx->stats.n_grouped * 48;
    return done; // the caller waits on it and hands it back to ctx->evpool
}

struct HostWriteback {
    std::thread worker;
    DBuf<double> out;                 // 7 words per grouped particle: index, pos, vel -- alive until the join
    cudaEvent_t ready = nullptr, done[2] = {nullptr, nullptr};
    cudaError_t err = cudaSuccess;
    size_t bytes = 0;
};

HostWriteback *hostwb_start(vrgpu_ctx *ctx, const vrgpu_view &view, const DeviceParticles &P, const long long *pfof_dev)
{
    if (P.n == 0 || mapped_host_pointer(view.base)) return nullptr;
    VR_REQUIRE(P.has_vel && view.off_vel != VRGPU_NOFIELD, VRGPU_EINVAL, "write-back needs velocities in the view");
    const int T = 256;
    uint32_t cnt = 0;
    std::unique_ptr<HostWriteback> job(new HostWriteback);
    {
        DBuf<uint32_t> flags(ctx, P.n), excl(ctx, P.n), total(ctx, 1);
        VR_LAUNCH(ctx, k_flag_pos, grid_for(P.n, T), T, 0, pfof_dev, P.n, flags.p);
        exclusive_scan_u32(ctx, flags.p, excl.p, P.n, total.p);
        VR_CUDA(cudaMemcpyAsync(&cnt, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        if (cnt == 0) return nullptr;
        job->out.alloc(ctx, 7 * (size_t)cnt);
        VR_LAUNCH(ctx, k_pack_phase, grid_for(P.n, T), T, 0, P.pos.p, P.vel.p, flags.p, excl.p, P.n, job->out.p);
    }
    if (!ctx->copy_stream) VR_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    if (!ctx->stage_host[0]) {
        VR_CUDA(cudaHostAlloc(&ctx->stage_host[0], (size_t)32 << 20, cudaHostAllocDefault));
        VR_CUDA(cudaHostAlloc(&ctx->stage_host[1], (size_t)32 << 20, cudaHostAllocDefault));
    }
    job->ready = KScope::get(ctx); job->done[0] = KScope::get(ctx); job->done[1] = KScope::get(ctx);
    job->bytes = (size_t)cnt * 56;
    VR_CUDA(cudaEventRecord(job->ready, ctx->stream));
    VR_CUDA(cudaStreamWaitEvent(ctx->copy_stream, job->ready, 0));
    // everything the worker touches, by value: it never looks at the context
    HostWriteback *j = job.get();
    const int device = ctx->device;
    cudaStream_t cs = ctx->copy_stream;
    void *stage[2] = {ctx->stage_host[0], ctx->stage_host[1]};
    unsigned char *base = const_cast<unsigned char *>(static_cast<const unsigned char *>(view.base));
    const size_t stride = view.stride, op = view.off_pos, ov = view.off_vel;
    job->worker = std::thread([j, device, cs, stage, base, stride, op, ov] {
        void *st[2] = {stage[0], stage[1]};
        cudaError_t e = cudaSetDevice(device);
        size_t chunk = (size_t)32 << 20;
        chunk -= chunk % 56;
        const unsigned char *src = reinterpret_cast<const unsigned char *>(j->out.p);
        const size_t nchunks = (j->bytes + chunk - 1) / chunk;
        auto issue = [&](size_t c) {
            const size_t o = c * chunk, len = std::min(chunk, j->bytes - o);
            if (e == cudaSuccess) e = cudaMemcpyAsync(st[c & 1], src + o, len, cudaMemcpyDeviceToHost, cs);
            if (e == cudaSuccess) e = cudaEventRecord(j->done[c & 1], cs);
        };
        auto scatter = [&](const unsigned char *from, size_t len, int part, int nparts) {
            const size_t nrec = len / 56, k0 = nrec * part / nparts, k1 = nrec * (part + 1) / nparts;
            for (size_t k = k0; k < k1; k++) {
                const double *r = reinterpret_cast<const double *>(from + 56 * k);
                unsigned long long i;
                memcpy(&i, r, 8);
                unsigned char *rec = base + (size_t)i * stride;
                double *pp = reinterpret_cast<double *>(rec + op);
                double *vv = reinterpret_cast<double *>(rec + ov);
                pp[0] = r[1]; pp[1] = r[2]; pp[2] = r[3];
                vv[0] = r[4]; vv[1] = r[5]; vv[2] = r[6];
            }
        };
        issue(0);
        for (size_t c = 0; c < nchunks && e == cudaSuccess; c++) {
            e = cudaEventSynchronize(j->done[c & 1]);
            if (e != cudaSuccess) break;
            if (c + 1 < nchunks) issue(c + 1);
            const size_t len = std::min(chunk, j->bytes - c * chunk);
            const unsigned char *from = static_cast<const unsigned char *>(st[c & 1]);
            const int nt = len > ((size_t)1 << 20) ? 8 : 1;
            std::thread th[8];
            for (int t = 1; t < nt; t++) th[t] = std::thread([&, t] { scatter(from, len, t, nt); });
            scatter(from, len, 0, nt);
            for (int t = 1; t < nt; t++) th[t].join();
        }
        if (e != cudaSuccess) cudaStreamSynchronize(cs); // leave nothing in flight on the staging buffers
        j->err = e;
    });
    return job.release();
}

void hostwb_join(vrgpu_ctx *ctx, HostWriteback *job, bool may_throw)
{
    if (!job) return;
    std::unique_ptr<HostWriteback> own(job);
    if (job->worker.joinable()) job->worker.join();
    ctx->evpool.push_back(job->ready);
    ctx->evpool.push_back(job->done[0]);
    ctx->evpool.push_back(job->done[1]);
    ctx->stats.d2h_bytes += (long long)job->bytes;
    const cudaError_t e = job->err;
    own.reset(); // (the packed records go back to the pool: the copy stream is done with them)
    if (e != cudaSuccess && may_throw) VR_CUDA(e);
}

void writeback_phase(vrgpu_ctx *ctx, const vrgpu_view &view, const DeviceParticles &P, const long long *pfof_dev)
{
    if (P.n == 0) return;
    VR_REQUIRE(P.has_vel && view.off_vel != VRGPU_NOFIELD, VRGPU_EINVAL, "write-back needs velocities in the view");
    const int T = 256;
    if (void *dev = mapped_host_pointer(view.base)) {
        KScope kwb(ctx, "io.writeback");
        VR_LAUNCH(ctx, k_writeback_direct, grid_for(6 * P.n, T), T, 0, static_cast<unsigned char *>(dev), view.stride, view.off_pos, view.off_vel,
                  P.pos.p, P.vel.p, pfof_dev, P.n);
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->stats.d2h_bytes += (long long)ctx->stats.n_grouped * 48;
        return;
    }
    DBuf<uint32_t> flags(ctx, P.n), excl(ctx, P.n), total(ctx, 1);
    VR_LAUNCH(ctx, k_flag_pos, grid_for(P.n, T), T, 0, pfof_dev, P.n, flags.p);
    exclusive_scan_u32(ctx, flags.p, excl.p, P.n, total.p);
    uint32_t cnt = 0;
    VR_CUDA(cudaMemcpyAsync(&cnt, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    if (cnt == 0) return;
    // pageable caller memory: 56-byte records (index, pos, vel) come back through the page-locked staging buffers and
    // host threads scatter them into the caller's records
    DBuf<double> out(ctx, 7 * (size_t)cnt);
    VR_LAUNCH(ctx, k_pack_phase, grid_for(P.n, T), T, 0, P.pos.p, P.vel.p, flags.p, excl.p, P.n, out.p);
    unsigned char *base = const_cast<unsigned char *>(static_cast<const unsigned char *>(view.base));
    const size_t stride = view.stride, op = view.off_pos, ov = view.off_vel;
    staged_download(ctx, out.p, (size_t)cnt * 56, 56, [=](const unsigned char *chunk, size_t, size_t len, int part, int nparts) {
        const size_t nrec = len / 56, k0 = nrec * part / nparts, k1 = nrec * (part + 1) / nparts;
        for (size_t k = k0; k < k1; k++) {
            const double *r = reinterpret_cast<const double *>(chunk + 56 * k);
            unsigned long long i;
            memcpy(&i, r, 8);
            unsigned char *rec = base + (size_t)i * stride;
            double *p = reinterpret_cast<double *>(rec + op);
            double *v = reinterpret_cast<double *>(rec + ov);
            p[0] = r[1]; p[1] = r[2]; p[2] = r[3];
            v[0] = r[4]; v[1] = r[5]; v[2] = r[6];
        }
    });
}

} // namespace vrgpu
