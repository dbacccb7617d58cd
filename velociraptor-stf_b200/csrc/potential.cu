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

namespace vrgpu {

constexpr int PT = 256; // threads per CTA
constexpr int PI = 4;   // slots per thread (register tiling: every shared-memory j serves PI pair evaluations)
constexpr int PS = PT * PI;

// 1/sqrt(r2): hardware seed (rsqrt.approx.f64 = MUFU.RSQ64H on the high word, ~2^-22) + one fp64 Newton step (4 DP
// instructions, relative error ~1e-13); zero / denormal / huge arguments take the exact path (inf for r2 == 0, like the
// reference).
__device__ __forceinline__ double rsqrt_refined(double r2)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(r2));
    if (!(r2 > 1e-290) || !(r2 < 1e290)) return 1.0 / sqrt(r2);
    const double h = 0.5 * r2;
    y = y * (1.5 - h * y * y);
    return y;
}

__global__ void __launch_bounds__(PT) k_potential_direct(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                                                         const double *__restrict__ m, const uint32_t *__restrict__ gid,
                                                         const uint32_t *__restrict__ off, uint32_t ntot, double G, double eps2,
                                                         long long pp_threshold, double *__restrict__ phi)
{
    __shared__ double2 sxy[PT], szm[PT]; // 16-byte shared loads: two per j instead of four
    __shared__ uint32_t sg[PT];
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
            mine[k] = pp_threshold < 0 || (long long)sz_g <= pp_threshold;
        }
        any_mine = any_mine || mine[k];
    }
    if (!__syncthreads_or(any_mine)) return;
    const uint32_t g_first = gid[s_first], g_last = gid[s_last];
    const uint32_t j0 = off[g_first], j1 = off[g_last + 1];
    const bool uniform = g_first == g_last; // the whole CTA sits inside one group (implies every slot is `mine`)
    for (uint32_t jb = j0; jb < j1; jb += PT) {
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
#pragma unroll
                for (int k = 0; k < PI; k++) {
                    const double dx = xi[k] - qx, dy = yi[k] - qy, dz = zi[k] - qz;
                    const double r2 = fma(dz, dz, fma(dy, dy, fma(dx, dx, eps2)));
                    acc[k] = fma(qm, rsqrt_refined(r2), acc[k]);
                }
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

unsigned long long potential_direct(vrgpu_ctx *ctx, const GroupSet &gs, double G, double eps, long long pp_threshold, double *phi)
{
    if (gs.ntot == 0) return 0;
    DBuf<unsigned long long> np(ctx, 1);
    VR_CUDA(cudaMemsetAsync(np.p, 0, sizeof(unsigned long long), ctx->stream));
    VR_LAUNCH(ctx, k_count_pairs, grid_for((size_t)gs.ngroups, 256), 256, 0, gs.off.p, (uint32_t)gs.ngroups, pp_threshold, np.p);
    VR_LAUNCH(ctx, k_potential_direct, grid_for(gs.ntot, PS), PT, 0, gs.x.p, gs.y.p, gs.z.p, gs.m.p, gs.gid.p, gs.off.p, (uint32_t)gs.ntot, G,
              eps * eps, pp_threshold, phi);
    unsigned long long h = 0;
    VR_CUDA(cudaMemcpyAsync(&h, np.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return h;
}

} // namespace vrgpu
