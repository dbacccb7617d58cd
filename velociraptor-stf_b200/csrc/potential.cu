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

constexpr int PT = 256;

__device__ __forceinline__ double rsqrt_refined(double r2)
{
    float f = (float)r2;
    double y = (double)rsqrtf(f);
    if (!(f > 0.0f) || !(f < 3.0e38f)) return 1.0 / sqrt(r2); // zero / denormal / huge: exact path (inf for r2 == 0, like the reference)
    double h = 0.5 * r2;
    y = y * (1.5 - h * y * y);
    return y;
}

__global__ void __launch_bounds__(PT) k_potential_direct(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                                                         const double *__restrict__ m, const uint32_t *__restrict__ gid,
                                                         const uint32_t *__restrict__ off, uint32_t ntot, double G, double eps2,
                                                         long long pp_threshold, double *__restrict__ phi)
{
    __shared__ double sx[PT], sy[PT], sz[PT], sm[PT];
    __shared__ uint32_t sg[PT];
    const uint32_t s_first = blockIdx.x * PT;
    const uint32_t s = s_first + threadIdx.x;
    const uint32_t s_last = min(ntot, s_first + PT) - 1;
    const bool active = s < ntot;
    uint32_t g = 0;
    double xi = 0, yi = 0, zi = 0, mi = 0;
    bool mine = false;
    if (active) {
        g = gid[s];
        xi = x[s]; yi = y[s]; zi = z[s]; mi = m[s];
        uint32_t sz_g = off[g + 1] - off[g];
        mine = pp_threshold < 0 || (long long)sz_g <= pp_threshold;
    }
    if (!__syncthreads_or(mine)) return;
    const uint32_t g_first = gid[s_first], g_last = gid[s_last];
    const uint32_t j0 = off[g_first], j1 = off[g_last + 1];
    const bool uniform = g_first == g_last;
    double acc = 0.0;
    for (uint32_t jb = j0; jb < j1; jb += PT) {
        uint32_t j = jb + threadIdx.x;
        if (j < j1) { sx[threadIdx.x] = x[j]; sy[threadIdx.x] = y[j]; sz[threadIdx.x] = z[j]; sm[threadIdx.x] = m[j]; sg[threadIdx.x] = gid[j]; }
        else { sg[threadIdx.x] = 0; sm[threadIdx.x] = 0; sx[threadIdx.x] = 0; sy[threadIdx.x] = 0; sz[threadIdx.x] = 0; }
        __syncthreads();
        if (mine) {
            const int cnt = (int)min((uint32_t)PT, j1 - jb);
            if (uniform) {
#pragma unroll 4
                for (int jj = 0; jj < cnt; jj++) {
                    double dx = xi - sx[jj], dy = yi - sy[jj], dz = zi - sz[jj];
                    double r2 = dx * dx + dy * dy + dz * dz + eps2;
                    double w = (jb + jj != s) ? sm[jj] : 0.0;
                    double ri = (jb + jj != s) ? rsqrt_refined(r2) : 0.0;
                    acc += w * ri;
                }
            } else {
                for (int jj = 0; jj < cnt; jj++) {
                    if (sg[jj] != g || jb + jj == s) continue;
                    double dx = xi - sx[jj], dy = yi - sy[jj], dz = zi - sz[jj];
                    double r2 = dx * dx + dy * dy + dz * dz + eps2;
                    acc += sm[jj] * rsqrt_refined(r2);
                }
            }
        }
        __syncthreads();
    }
    if (mine) phi[s] = -G * mi * acc;
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
    VR_LAUNCH(ctx, k_potential_direct, grid_for(gs.ntot, PT), PT, 0, gs.x.p, gs.y.p, gs.z.p, gs.m.p, gs.gid.p, gs.off.p, (uint32_t)gs.ntot, G,
              eps * eps, pp_threshold, phi);
    unsigned long long h = 0;
    VR_CUDA(cudaMemcpyAsync(&h, np.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return h;
}

} // namespace vrgpu
