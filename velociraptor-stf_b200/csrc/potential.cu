// velociraptor-stf_b200/csrc/potential.cu -- per-group gravitational potentials, direct sum.
//
// Replaces the pair loop of Unbind (src/unbind.cxx:398-416): phi_j = -G * sum_{k != j} m_j m_k / sqrt(r_jk^2 + eps^2),
// fp64 throughout (positions are differenced in fp64; the reciprocal square root is an fp32 seed refined by one
// fp64 Newton step, relative error ~1e-13, far inside the 1e-6 tolerance of BASELINE.json).
//
// Layout: the group-major SoA of groups.cuh.  One CTA owns 256 consecutive SLOTS (they may span several small
// groups or be a slice of one big group); the j-range it has to visit is [off[g_first], off[g_last + 1]) and is
// streamed through shared memory in tiles of 256.  FP64-pipe bound: ~14 DP instructions per pair.
#include "groups.cuh"
#include "rsqrt.cuh"

#include <algorithm>
#include <cstdlib>

namespace vrgpu {

constexpr int PT = 256; // threads per CTA
constexpr int PI = 4;   // slots per thread (register tiling: every shared-memory j serves PI pair evaluations)
constexpr int PS = PT * PI;

__global__ void __launch_bounds__(PT) k_potential_direct(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                                                         const double *__restrict__ m, const uint32_t *__restrict__ gid,
                                                         const uint32_t *__restrict__ off, uint32_t ntot, double G, double eps2,
                                                         long long pp_threshold, uint32_t sym_min, double *__restrict__ phi)
{
    __shared__ double2 sxy[PT], szm[PT]; // 16-byte shared loads: two per j instead of four
    __shared__ uint32_t sg[PT];
    __shared__ uint32_t s_lo, s_hi;
    const uint32_t s_first = blockIdx.x * PS;
    const uint32_t s_last = min(ntot, s_first + PS) - 1;
    uint32_t s[PI], g[PI];
    double xi[PI], yi[PI], zi[PI], mi[PI], acc[PI];
    bool mine[PI];
    bool any_mine = false;
#pragma unroll
    for (int k = 0; k < PI; k++) {
        s[k] = s_first + k * PT + threadIdx.x;
        g[k] = 0; xi[k] = yi[k] = zi[k] = mi[k] = 0; acc[k] = 0; mine[k] = false;
        if (s[k] < ntot) {
            g[k] = gid[s[k]];
            xi[k] = x[s[k]]; yi[k] = y[s[k]]; zi[k] = z[s[k]]; mi[k] = m[s[k]];
            const uint32_t sz_g = off[g[k] + 1] - off[g[k]];
            mine[k] = (pp_threshold < 0 || (long long)sz_g <= pp_threshold) && sz_g < sym_min;
        }
        any_mine = any_mine || mine[k];
    }
    // the j range runs from the first to the last group this CTA actually serves: a CTA that holds the tail of a group
    // left to the symmetric kernel must not walk that (possibly huge) group
    if (threadIdx.x == 0) { s_lo = 0xffffffffu; s_hi = 0u; }
    if (!__syncthreads_or(any_mine)) return;
#pragma unroll
    for (int k = 0; k < PI; k++)
        if (mine[k]) { atomicMin(&s_lo, s[k]); atomicMax(&s_hi, s[k]); }
    __syncthreads();
    const uint32_t g_first = gid[s_lo], g_last = gid[s_hi];
    const uint32_t j0 = off[g_first], j1 = off[g_last + 1];
    const bool uniform = gid[s_first] == gid[s_last]; // the whole CTA sits inside one group (implies every slot is `mine`)
    for (uint32_t jb = j0; jb < j1; jb += PT) {
        // a tile that lies wholly inside a group this kernel does not serve (ids in arbitrary order) is skipped
        {
            const uint32_t ga = gid[jb];
            const uint32_t ea = off[ga + 1], na = ea - off[ga];
            const bool served = (pp_threshold < 0 || (long long)na <= pp_threshold) && na < sym_min;
            if (!served && ea >= jb + PT) {
                jb = j0 + ((ea - j0) / PT) * PT - PT; // the loop increment then lands on the tile that holds the group's end
                continue;
            }
        }
        const uint32_t j = jb + threadIdx.x;
        if (j < j1) { sxy[threadIdx.x] = make_double2(x[j], y[j]); szm[threadIdx.x] = make_double2(z[j], m[j]); sg[threadIdx.x] = gid[j]; }
        else { sg[threadIdx.x] = 0; sxy[threadIdx.x] = make_double2(0, 0); szm[threadIdx.x] = make_double2(0, 0); }
        __syncthreads();
        const int cnt = (int)min((uint32_t)PT, j1 - jb);
        // a tile that holds none of this CTA's own slots needs no self-exclusion test at all
        const bool has_self = jb < s_first + PS && jb + PT > s_first;
        if (uniform && !has_self) {
#pragma unroll 4
            for (int jj = 0; jj < cnt; jj++) {
                const double2 qa = sxy[jj], qb = szm[jj];
                const double qx = qa.x, qy = qa.y, qz = qb.x, qm = qb.y;
                double r2[PI], ri[PI];
                unsigned worst = 0;
#pragma unroll
                for (int k = 0; k < PI; k++) {
                    const double dx = xi[k] - qx, dy = yi[k] - qy, dz = zi[k] - qz;
                    r2[k] = fma(dz, dz, fma(dy, dy, fma(dx, dx, eps2)));
                    const double y0 = rsqrt_seed(r2[k]);
                    worst = max(worst, (unsigned)__double2hiint(y0));
                    ri[k] = rsqrt_newton(r2[k], y0);
                }
                if (worst >= 0x5f300000u) {
#pragma unroll
                    for (int k = 0; k < PI; k++) ri[k] = rsqrt_refined(r2[k]);
                }
#pragma unroll
                for (int k = 0; k < PI; k++) acc[k] = fma(qm, ri[k], acc[k]);
            }
        } else {
            for (int jj = 0; jj < cnt; jj++) {
                const double2 qa = sxy[jj], qb = szm[jj];
                const double qx = qa.x, qy = qa.y, qz = qb.x, qm = qb.y;
                const uint32_t qg = sg[jj];
#pragma unroll
                for (int k = 0; k < PI; k++) {
                    if (!mine[k] || qg != g[k] || jb + jj == s[k]) continue;
                    const double dx = xi[k] - qx, dy = yi[k] - qy, dz = zi[k] - qz;
                    const double r2 = fma(dz, dz, fma(dy, dy, fma(dx, dx, eps2)));
                    acc[k] = fma(qm, rsqrt_refined(r2), acc[k]);
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < PI; k++)
        if (mine[k]) phi[s[k]] = -G * mi[k] * acc[k];
}

// ---------------------------------------------------------------------------------------------------
// symmetric direct sum for big groups: every unordered TILE pair (I <= J) of a group is one CTA.  The 1/r of a particle
// pair is evaluated once and credited to both particles: acc_i in registers, the j side accumulated per warp in shared memory
// and over the CTA through shared memory (lanes rotate over the j of a 32-block, so no shuffles are needed).  45 FP64-pipe
// instructions per 4 unordered pairs (3 DADD + 3 DFMA + 3 Newton + 2 accumulate each, + 1 for the j side).
// Partial sums of different CTAs meet in global fp64 atomics, so the last bits of phi depend on the arrival order
// (relative 1e-16; DESIGN.md "determinism").
// ---------------------------------------------------------------------------------------------------
constexpr int SYM_PT = 128, SYM_PI = 8;    // threads per CTA, i-slots per thread of the symmetric kernel
constexpr int SYM_TILE = SYM_PT * SYM_PI; // 1024 slots
constexpr uint32_t SYM_MIN = 4 * SYM_TILE; // groups at least this large take the symmetric kernel

__global__ void k_sym_count(const uint32_t *__restrict__ off, uint32_t ngroups, long long pp_threshold, unsigned long long *__restrict__ npairs)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; // entry g holds group g (entry 0 unused)
    if (g > ngroups) return;
    unsigned long long c = 0;
    if (g >= 1) {
        const uint32_t n = off[g + 1] - off[g];
        if (n >= SYM_MIN && (pp_threshold < 0 || (long long)n <= pp_threshold)) {
            const unsigned long long nt = (n + SYM_TILE - 1) / SYM_TILE;
            c = nt * (nt + 1) / 2;
        }
    }
    npairs[g] = c;
}

// 128 threads x 8 slots, 3 CTAs per SM (168 registers) and no unrolling of the rotation loop measured best: more slots per
// thread amortise the shared-memory traffic and loop overhead of a rotation step over more FP64 work (256 x 4: 106 ms,
// 128 x 8: 99.5 ms at 256^3)
// Measured and dropped (profiles/r02a_probe.md): a j-side sum that rotates through the lanes with a register shuffle
// instead of the shared read-modify-write, with and without a prefetch of the next step's j -- both 2-3 % slower.
template <int SPT, int SPI, int MINB>
__global__ void __launch_bounds__(SPT, MINB) k_potential_sym(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                                                      const double *__restrict__ m, const uint32_t *__restrict__ off, uint32_t ngroups,
                                                      const unsigned long long *__restrict__ pair_start /* exclusive scan, ngroups+2 */,
                                                      double eps2, double *__restrict__ phiacc)
{
    static_assert(SPT * SPI == SYM_TILE, "tile size is fixed");
    __shared__ double2 sxy[SPT], szm[SPT];
    __shared__ double sacc[SPT / 32][SPT];
    __shared__ uint32_t s_g;
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    // which group: largest g with pair_start[g] <= blockIdx.x
    if (threadIdx.x == 0) {
        uint32_t lo = 1, hi = ngroups;
        const unsigned long long b = blockIdx.x;
        while (lo < hi) {
            const uint32_t mid = (lo + hi + 1) >> 1;
            if (pair_start[mid] <= b) lo = mid; else hi = mid - 1;
        }
        s_g = lo;
    }
    __syncthreads();
    const uint32_t g = s_g;
    const uint32_t g0 = off[g], n = off[g + 1] - g0;
    const uint32_t nt = (n + SYM_TILE - 1) / SYM_TILE;
    // triangular decode: t = I*nt - I*(I-1)/2 + (J - I)
    const unsigned long long t = blockIdx.x - pair_start[g];
    uint32_t I = (uint32_t)floor(((2.0 * nt + 1.0) - sqrt((2.0 * nt + 1.0) * (2.0 * nt + 1.0) - 8.0 * (double)t)) * 0.5);
    while ((unsigned long long)I * nt - (unsigned long long)I * (I - 1) / 2 > t) I--;
    while ((unsigned long long)(I + 1) * nt - (unsigned long long)(I + 1) * I / 2 <= t) I++;
    const uint32_t J = I + (uint32_t)(t - ((unsigned long long)I * nt - (unsigned long long)I * (I - 1) / 2));
    const uint32_t i_base = g0 + I * SYM_TILE, j_base = g0 + J * SYM_TILE, g_end = g0 + n;
    const uint32_t j_end = min(g_end, j_base + SYM_TILE);
    const bool diag = I == J;

    uint32_t si[SPI];
    double xi[SPI], yi[SPI], zi[SPI], mi[SPI], acc[SPI];
#pragma unroll
    for (int k = 0; k < SPI; k++) {
        si[k] = i_base + k * SPT + threadIdx.x;
        acc[k] = 0;
        if (si[k] < g_end) { xi[k] = x[si[k]]; yi[k] = y[si[k]]; zi[k] = z[si[k]]; mi[k] = m[si[k]]; }
        else { xi[k] = 1e100; yi[k] = 1e100; zi[k] = 1e100; mi[k] = 0; } // padding: weightless and far away
    }
    for (uint32_t jb = j_base; jb < j_end; jb += SPT) {
        const uint32_t j = jb + threadIdx.x;
        if (j < j_end) { sxy[threadIdx.x] = make_double2(x[j], y[j]); szm[threadIdx.x] = make_double2(z[j], m[j]); }
        else { sxy[threadIdx.x] = make_double2(-1e100, -1e100); szm[threadIdx.x] = make_double2(-1e100, 0.0); }
#pragma unroll
        for (int ww = 0; ww < SPT / 32; ww++) sacc[ww][threadIdx.x] = 0.0;
        __syncthreads();
        const int cnt = (int)min((uint32_t)SPT, j_end - jb);
        if (!diag) {
            // rotation: at step s lane l works on j = 32*blk + ((l + s) & 31), so the 32 lanes always hold 32 DIFFERENT j:
            // the j-side partial goes straight into the warp's shared row without any cross-lane reduction
            const int nblk = (cnt + 31) >> 5;
            for (int blk = 0; blk < nblk; blk++) {
#pragma unroll 1
                for (int st = 0; st < 32; st++) {
                    const int jj = (blk << 5) + (int)((lane + st) & 31);
                    const double2 qa = sxy[jj], qb = szm[jj];
                    double tj = 0, r2[SPI], ri[SPI];
                    unsigned worst = 0;
#pragma unroll
                    for (int k = 0; k < SPI; k++) {
                        const double dx = xi[k] - qa.x, dy = yi[k] - qa.y, dz = zi[k] - qb.x;
                        r2[k] = fma(dz, dz, fma(dy, dy, fma(dx, dx, eps2)));
                        const double y0 = rsqrt_seed(r2[k]);
                        worst = max(worst, (unsigned)__double2hiint(y0));
                        ri[k] = rsqrt_newton(r2[k], y0);
                    }
                    if (worst >= 0x5f300000u) {
#pragma unroll
                        for (int k = 0; k < SPI; k++) ri[k] = rsqrt_refined(r2[k]);
                    }
#pragma unroll
                    for (int k = 0; k < SPI; k++) {
                        acc[k] = fma(qb.y, ri[k], acc[k]);
                        tj = fma(mi[k], ri[k], tj);
                    }
                    sacc[w][jj] += tj;
                    __syncwarp();
                }
            }
            __syncthreads();
            if ((int)threadIdx.x < cnt) {
                double s = 0;
#pragma unroll
                for (int ww = 0; ww < SPT / 32; ww++) s += sacc[ww][threadIdx.x];
                atomicAdd(&phiacc[jb + threadIdx.x], s);
            }
        } else {
            for (int jj = 0; jj < cnt; jj++) {
                const double2 qa = sxy[jj], qb = szm[jj];
                const uint32_t sj = jb + jj;
#pragma unroll
                for (int k = 0; k < SPI; k++) {
                    if (sj == si[k]) continue;
                    const double dx = xi[k] - qa.x, dy = yi[k] - qa.y, dz = zi[k] - qb.x;
                    const double r2 = fma(dz, dz, fma(dy, dy, fma(dx, dx, eps2)));
                    acc[k] = fma(qb.y, rsqrt_refined(r2), acc[k]);
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < SPI; k++)
        if (si[k] < g_end) atomicAdd(&phiacc[si[k]], acc[k]);
}

__global__ void k_sym_finish(const double *__restrict__ phiacc, const double *__restrict__ m, const uint32_t *__restrict__ gid,
                             const uint32_t *__restrict__ off, uint32_t ntot, double G, long long pp_threshold, double *__restrict__ phi)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot) return;
    const uint32_t g = gid[s];
    const uint32_t n = off[g + 1] - off[g];
    if (n >= SYM_MIN && (pp_threshold < 0 || (long long)n <= pp_threshold)) phi[s] = -G * m[s] * phiacc[s];
}

__global__ void k_count_pairs(const uint32_t *__restrict__ off, uint32_t ngroups, long long pp_threshold, unsigned long long *__restrict__ out)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x + 1;
    unsigned long long c = 0;
    if (g <= ngroups) {
        unsigned long long n = off[g + 1] - off[g];
        if (pp_threshold < 0 || (long long)n <= pp_threshold) c = n * (n ? n - 1 : 0);
    }
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

unsigned long long potential_direct(vrgpu_ctx *ctx, const GroupSet &gs, double G_, double eps, long long pp_threshold, double *phi)
{
    if (gs.ntot == 0) return 0;
    DBuf<unsigned long long> np(ctx, 1);
    VR_CUDA(cudaMemsetAsync(np.p, 0, sizeof(unsigned long long), ctx->stream));
    VR_LAUNCH(ctx, k_count_pairs, grid_for((size_t)gs.ngroups, 256), 256, 0, gs.off.p, (uint32_t)gs.ngroups, pp_threshold, np.p);
    // big groups: symmetric tile-pair kernel; everything else: the tiled kernel
    const uint32_t G = (uint32_t)gs.ngroups;
    DBuf<unsigned long long> npairs(ctx, (size_t)G + 2), pstart(ctx, (size_t)G + 2), ptotal(ctx, 1);
    VR_CUDA(cudaMemsetAsync(npairs.p, 0, sizeof(unsigned long long) * ((size_t)G + 2), ctx->stream));
    VR_LAUNCH(ctx, k_sym_count, grid_for((size_t)G + 1, 256), 256, 0, gs.off.p, G, pp_threshold, npairs.p);
    exclusive_scan_u64(ctx, npairs.p, pstart.p, (size_t)G + 2, ptotal.p);
    unsigned long long nwork = 0;
    VR_CUDA(cudaMemcpyAsync(&nwork, ptotal.p, sizeof(nwork), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    VR_REQUIRE(nwork < 0x7fffffffull, VRGPU_EUNSUPPORTED, "direct sum: too many tile pairs (use the tree for groups this large)");
    { KScope k(ctx, "potential.direct.tiled");
    VR_LAUNCH(ctx, k_potential_direct, grid_for(gs.ntot, PS), PT, 0, gs.x.p, gs.y.p, gs.z.p, gs.m.p, gs.gid.p, gs.off.p, (uint32_t)gs.ntot, G_,
              eps * eps, pp_threshold, nwork ? SYM_MIN : 0xffffffffu, phi); }
    if (nwork) {
        KScope k(ctx, "potential.direct.sym");
        DBuf<double> phiacc(ctx, gs.ntot);
        VR_CUDA(cudaMemsetAsync(phiacc.p, 0, sizeof(double) * gs.ntot, ctx->stream));
        VR_LAUNCH(ctx, (k_potential_sym<SYM_PT, SYM_PI, 3>), (unsigned)nwork, SYM_PT, 0, gs.x.p, gs.y.p, gs.z.p, gs.m.p, gs.off.p, G, pstart.p,
                  eps * eps, phiacc.p);
        VR_LAUNCH(ctx, k_sym_finish, grid_for(gs.ntot, 256), 256, 0, phiacc.p, gs.m.p, gs.gid.p, gs.off.p, (uint32_t)gs.ntot, G_, pp_threshold, phi);
    }
    unsigned long long h = 0;
    VR_CUDA(cudaMemcpyAsync(&h, np.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return h;
}

} // namespace vrgpu

// ---------------------------------------------------------------------------------------------------
// FP64 pipe microbenchmark (roofline denominator for the potential kernels; MEASURED_PEAKS.json has no fp64 figure)
// ---------------------------------------------------------------------------------------------------
namespace vrgpu {
__global__ void __launch_bounds__(256) k_dfma_peak(double *out, int iters, double a, double b)
{
    double r0 = threadIdx.x, r1 = r0 + 1, r2 = r0 + 2, r3 = r0 + 3, r4 = r0 + 4, r5 = r0 + 5, r6 = r0 + 6, r7 = r0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            r0 = fma(r0, a, b); r1 = fma(r1, a, b); r2 = fma(r2, a, b); r3 = fma(r3, a, b);
            r4 = fma(r4, a, b); r5 = fma(r5, a, b); r6 = fma(r6, a, b); r7 = fma(r7, a, b);
        }
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r0 + r1 + r2 + r3 + r4 + r5 + r6 + r7;
}

double measure_fp64_tflops(vrgpu_ctx *ctx)
{
    const int blocks = ctx->sm_count * 8, iters = 2048;
    DBuf<double> out(ctx, (size_t)blocks * 256);
    VR_LAUNCH(ctx, k_dfma_peak, blocks, 256, 0, out.p, 64, 0.999999, 1e-9); // warm-up
    cudaEvent_t a = KScope::get(ctx), b = KScope::get(ctx);
    double best = 0;
    for (int rep = 0; rep < 3; rep++) {
        VR_CUDA(cudaEventRecord(a, ctx->stream));
        VR_LAUNCH(ctx, k_dfma_peak, blocks, 256, 0, out.p, iters, 0.999999, 1e-9);
        VR_CUDA(cudaEventRecord(b, ctx->stream));
        VR_CUDA(cudaEventSynchronize(b));
        float ms = 0;
        VR_CUDA(cudaEventElapsedTime(&ms, a, b));
        const double flops = 2.0 * 8 * 16 * (double)iters * 256.0 * blocks;
        best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    ctx->evpool.push_back(a);
    ctx->evpool.push_back(b);
    return best;
}

// Probe for the next design step of the direct sum: does the fp64 tensor-core instruction
// (mma.sync m8n8k4 f64, 512 flops per warp instruction) run beside the vector DFMA pipe or on it?  If beside, the
// r^2 = |xi|^2 + |xj|^2 - 2 xi.xj part of a tile pair (6 of the 11 fp64 instructions per particle pair) can move to
// it.  mode 0: DFMA only, 1: DMMA only, 2: both in one instruction stream (same counts as 0 and 1 together).
template <int MODE> __global__ void __launch_bounds__(256) k_fp64_pipes(double *out, int iters, double a, double b)
{
    double r0 = threadIdx.x, r1 = r0 + 1, r2 = r0 + 2, r3 = r0 + 3, r4 = r0 + 4, r5 = r0 + 5, r6 = r0 + 6, r7 = r0 + 7;
    double c0 = r0, c1 = r1, c2 = r2, c3 = r3, c4 = r4, c5 = r5, c6 = r6, c7 = r7; // four independent DMMA accumulator pairs
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            if (MODE != 1) {
                r0 = fma(r0, a, b); r1 = fma(r1, a, b); r2 = fma(r2, a, b); r3 = fma(r3, a, b);
                r4 = fma(r4, a, b); r5 = fma(r5, a, b); r6 = fma(r6, a, b); r7 = fma(r7, a, b);
            }
            if (MODE != 0) {
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c2), "+d"(c3) : "d"(a), "d"(b));
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c4), "+d"(c5) : "d"(a), "d"(b));
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c6), "+d"(c7) : "d"(a), "d"(b));
            }
        }
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r0 + r1 + r2 + r3 + r4 + r5 + r6 + r7 + c0 + c1 + c2 + c3 + c4 + c5 + c6 + c7;
}

// out[0]: DFMA-only TFLOP/s, out[1]: DMMA-only TFLOP/s, out[2] / out[3]: DFMA / DMMA TFLOP/s when both are issued together
void probe_fp64_pipes(vrgpu_ctx *ctx, double out4[4])
{
    const int blocks = ctx->sm_count * 8, iters = 1024;
    DBuf<double> out(ctx, (size_t)blocks * 256);
    cudaEvent_t a = KScope::get(ctx), b = KScope::get(ctx);
    double ms_mode[3] = {0, 0, 0};
    for (int mode = 0; mode < 3; mode++) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; rep++) { // rep 0 is the warm-up
            VR_CUDA(cudaEventRecord(a, ctx->stream));
            if (mode == 0) VR_LAUNCH(ctx, k_fp64_pipes<0>, blocks, 256, 0, out.p, iters, 0.999999, 1e-9);
            else if (mode == 1) VR_LAUNCH(ctx, k_fp64_pipes<1>, blocks, 256, 0, out.p, iters, 0.999999, 1e-9);
            else VR_LAUNCH(ctx, k_fp64_pipes<2>, blocks, 256, 0, out.p, iters, 0.999999, 1e-9);
            VR_CUDA(cudaEventRecord(b, ctx->stream));
            VR_CUDA(cudaEventSynchronize(b));
            float ms = 0;
            VR_CUDA(cudaEventElapsedTime(&ms, a, b));
            if (rep) best = std::min(best, ms);
        }
        ms_mode[mode] = best;
    }
    ctx->evpool.push_back(a);
    ctx->evpool.push_back(b);
    const double dfma_flops = 2.0 * 8 * 16 * (double)iters * 256.0 * blocks;
    const double dmma_flops = 512.0 * 4 * 16 * (double)iters * 8.0 * blocks; // 512 flops per warp instruction, 8 warps per block
    out4[0] = dfma_flops / (ms_mode[0] * 1e-3) / 1e12;
    out4[1] = dmma_flops / (ms_mode[1] * 1e-3) / 1e12;
    out4[2] = dfma_flops / (ms_mode[2] * 1e-3) / 1e12;
    out4[3] = dmma_flops / (ms_mode[2] * 1e-3) / 1e12;
}
} // namespace vrgpu

// ---------------------------------------------------------------------------------------------------
// vrgpu_potential_direct_noalloc (include/vrgpu.h): the direct-sum potential of ONE group for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// It runs k_potential_direct -- the kernel vrgpu_potential / vrgpu_unbind use -- on a
// group-major SoA that a gather kernel writes into a FILE-SCOPE device array instead of the DBuf scratch of groups.cu, so
// the call allocates nothing (no cudaMalloc / cudaFree / stream-ordered pool, no context, no events).
//
// The array is 2.25 MiB and is held by every GPU that loads the library until exit.  All calls share it and it keeps the
// last call's contents: every word a call reads it first writes (columns [0, n), gid[0, n), off[0..2]), so nothing has to
// start at zero -- but two calls whose kernels overlap on the GPU (different streams, or different threads) RACE.
// ---------------------------------------------------------------------------------------------------
#include <cstdio>
#include <mutex>

namespace vrgpu {
constexpr long long NA_MAX = VRGPU_NOALLOC_MAX_PARTICLES;
static_assert(NA_MAX % 512 == 0, "column starts must keep the 256-byte alignment");
// layout in doubles: x | y | z | m (NA_MAX each) | gid (NA_MAX uint32 = NA_MAX/2 doubles) | off (4 uint32, padded to 32 doubles)
constexpr size_t NA_DOUBLES = (size_t)4 * NA_MAX + NA_MAX / 2 + 32;
static __device__ __align__(256) double g_noalloc_scratch[NA_DOUBLES];

// AoS positions (3 doubles per particle) and masses -> the SoA, group ids and offsets k_potential_direct reads: one group
// (id 1) in slots [0, n); off[0] = off[1] = 0, off[2] = n as in GroupSet (groups.cuh)
__global__ void k_noalloc_gather(const double *__restrict__ pos, const double *__restrict__ mass, uint32_t n, double *__restrict__ x,
                                 double *__restrict__ y, double *__restrict__ z, double *__restrict__ m, uint32_t *__restrict__ gid,
                                 uint32_t *__restrict__ off)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { off[0] = 0u; off[1] = 0u; off[2] = n; off[3] = n; }
    if (i >= n) return;
    x[i] = pos[3 * (size_t)i];
    y[i] = pos[3 * (size_t)i + 1];
    z[i] = pos[3 * (size_t)i + 2];
    m[i] = mass ? mass[i] : 1.0;
    gid[i] = 1u;
}

namespace {
constexpr int NA_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
double *g_noalloc_addr[NA_MAX_DEVICES]; // device address of g_noalloc_scratch on device d (the symbol itself is a host-side stand-in)

int noalloc_fail(char *errbuf, long long errbuf_len, int code, const char *what, const char *detail)
{
    if (errbuf && errbuf_len > 0) std::snprintf(errbuf, (size_t)errbuf_len, "vrgpu_potential_direct_noalloc: %s%s%s", what, detail ? ": " : "", detail ? detail : "");
    return code;
}
} // namespace
} // namespace vrgpu

extern "C" int vrgpu_potential_direct_noalloc(int device, void *stream, const double *pos_dev, const double *mass_dev, long long n, double G,
                                              double eps, double *phi_dev, char *errbuf, long long errbuf_len)
{
    using namespace vrgpu;
    if (errbuf && errbuf_len > 0) errbuf[0] = 0;
    if (n < 0) return noalloc_fail(errbuf, errbuf_len, VRGPU_EINVAL, "negative particle count", nullptr);
    if (n > NA_MAX) {
        char msg[160];
        std::snprintf(msg, sizeof msg, "%lld particles exceed the fixed scratch of VRGPU_NOALLOC_MAX_PARTICLES = %lld (use vrgpu_potential)", n, NA_MAX);
        return noalloc_fail(errbuf, errbuf_len, VRGPU_EINVAL, "group too large", msg);
    }
    if (n == 0) return VRGPU_OK;
    if (!pos_dev || !phi_dev) return noalloc_fail(errbuf, errbuf_len, VRGPU_EINVAL, "null device pointer (pos_dev / phi_dev)", nullptr);
    int cur = -1;
    cudaError_t e = cudaGetDevice(&cur);
    if (e != cudaSuccess) return noalloc_fail(errbuf, errbuf_len, VRGPU_ENODEV, "cudaGetDevice", cudaGetErrorString(e));
    if (device < 0) device = cur;
    if (device >= NA_MAX_DEVICES) return noalloc_fail(errbuf, errbuf_len, VRGPU_EINVAL, "device ordinal out of range", nullptr);
    if (device != cur) {
        e = cudaSetDevice(device);
        if (e != cudaSuccess) return noalloc_fail(errbuf, errbuf_len, VRGPU_ENODEV, "cudaSetDevice", cudaGetErrorString(e));
    }
    int rc = VRGPU_OK;
    double *base = nullptr;
    {
        // set-up, once per device, with that device current: the device address of the file-scope array
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_addr[device]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) rc = noalloc_fail(errbuf, errbuf_len, VRGPU_ECUDA, "cudaGetSymbolAddress", cudaGetErrorString(e));
            else g_noalloc_addr[device] = static_cast<double *>(p);
        }
        base = g_noalloc_addr[device];
    }
    if (rc == VRGPU_OK) {
        double *x = base, *y = base + NA_MAX, *z = base + 2 * NA_MAX, *m = base + 3 * NA_MAX;
        uint32_t *gid = reinterpret_cast<uint32_t *>(base + 4 * NA_MAX);
        uint32_t *off = reinterpret_cast<uint32_t *>(base + 4 * NA_MAX + NA_MAX / 2);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const uint32_t nn = (uint32_t)n;
        k_noalloc_gather<<<(nn + 255u) / 256u, 256, 0, st>>>(pos_dev, mass_dev, nn, x, y, z, m, gid, off);
        // pp_threshold < 0 and sym_min = 2^32 - 1: every slot is served by this kernel (no symmetric pass, no atomics)
        k_potential_direct<<<(nn + PS - 1) / PS, PT, 0, st>>>(x, y, z, m, gid, off, nn, G, eps * eps, -1ll, 0xffffffffu, phi_dev);
        e = cudaPeekAtLastError();
        if (e != cudaSuccess) rc = noalloc_fail(errbuf, errbuf_len, VRGPU_ECUDA, "kernel launch", cudaGetErrorString(e));
    }
    if (device != cur) cudaSetDevice(cur);
    return rc;
}
