// velociraptor-stf_b200/csrc/groups.cu -- group-major layout (see groups.cuh).
#include "groups.cuh"

#include <algorithm>

namespace vrgpu {

// Group ids come in size order from the search stages, so the low ids are the big groups: their counters are kept per CTA
// in shared memory (a hot global counter serialises ~10^5 atomics), every CTA walks a contiguous chunk of the array.
constexpr int PRIV_IDS = 4096, PRIV_THREADS = 512;

__global__ void __launch_bounds__(PRIV_THREADS) k_flag_grouped(const long long *__restrict__ pfof, size_t n, long long ngroups,
                                                                uint32_t *__restrict__ flags, uint32_t *__restrict__ count,
                                                                unsigned int *__restrict__ nbad)
{
    __shared__ uint32_t sc[PRIV_IDS];
    for (int k = threadIdx.x; k < PRIV_IDS; k += PRIV_THREADS) sc[k] = 0;
    __syncthreads();
    const size_t per = (n + gridDim.x - 1) / gridDim.x;
    const size_t i0 = (size_t)blockIdx.x * per, i1 = min(n, i0 + per);
    for (size_t i = i0 + threadIdx.x; i < i1; i += PRIV_THREADS) {
        long long g = pfof[i];
        if (g < 0 || g > ngroups) { atomicAdd(nbad, 1u); g = 0; }
        flags[i] = g > 0 ? 1u : 0u;
        if (g > 0) {
            if (g <= PRIV_IDS) atomicAdd(&sc[g - 1], 1u);
            else atomicAdd(&count[g], 1u);
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < PRIV_IDS; k += PRIV_THREADS)
        if (sc[k]) atomicAdd(&count[k + 1], sc[k]);
}

__global__ void k_compact_grouped(const long long *__restrict__ pfof, const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl,
                                  size_t n, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !flags[i]) return;
    uint32_t s = excl[i];
    keys[s] = (uint64_t)pfof[i];
    vals[s] = (uint32_t)i;
}

__global__ void k_keys_to_gid(const uint64_t *__restrict__ keys, size_t n, uint32_t *__restrict__ gid)
{
    size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) gid[s] = (uint32_t)keys[s];
}

__global__ void k_gather_cols(const double *__restrict__ pos, const double *__restrict__ vel, const double *__restrict__ mass,
                              const uint32_t *__restrict__ idx, size_t n, double *__restrict__ x, double *__restrict__ y, double *__restrict__ z,
                              double *__restrict__ vx, double *__restrict__ vy, double *__restrict__ vz, double *__restrict__ m)
{
    size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    size_t i = idx[s];
    x[s] = pos[3 * i]; y[s] = pos[3 * i + 1]; z[s] = pos[3 * i + 2];
    if (vx) { vx[s] = vel[3 * i]; vy[s] = vel[3 * i + 1]; vz[s] = vel[3 * i + 2]; }
    if (m) m[s] = mass ? mass[i] : 1.0;
}

static int bits_for_u64(uint64_t v)
{
    int b = 0;
    while (v) { b++; v >>= 1; }
    return b < 1 ? 1 : b;
}

void build_groupset(vrgpu_ctx *ctx, const long long *pfof_dev, size_t n, long long ngroups, GroupSet &gs)
{
    const int T = 256;
    gs.ngroups = ngroups;
    gs.ntot = 0;
    VR_REQUIRE(ngroups >= 0 && ngroups < 0x7fffffffLL, VRGPU_EINVAL, "bad group count");
    gs.off.alloc(ctx, (size_t)ngroups + 2);
    VR_CUDA(cudaMemsetAsync(gs.off.p, 0, sizeof(uint32_t) * ((size_t)ngroups + 2), ctx->stream));
    if (n == 0 || ngroups == 0) return;
    DBuf<uint32_t> flags(ctx, n), excl(ctx, n), count(ctx, (size_t)ngroups + 2), total(ctx, 1);
    DBuf<unsigned int> nbad(ctx, 1);
    VR_CUDA(cudaMemsetAsync(count.p, 0, sizeof(uint32_t) * ((size_t)ngroups + 2), ctx->stream));
    VR_CUDA(cudaMemsetAsync(nbad.p, 0, sizeof(unsigned int), ctx->stream));
    VR_LAUNCH(ctx, k_flag_grouped, (unsigned)std::max<size_t>(1, std::min<size_t>((size_t)ctx->sm_count * 4, (n + PRIV_THREADS - 1) / PRIV_THREADS)), PRIV_THREADS, 0,
              pfof_dev, n, ngroups, flags.p, count.p, nbad.p);
    exclusive_scan_u32(ctx, flags.p, excl.p, n, total.p);
    // off[g] = sum of count[0..g-1]; count[0] == 0 so off[0] = off[1] = 0
    exclusive_scan_u32(ctx, count.p, gs.off.p, (size_t)ngroups + 2, nullptr);
    uint32_t htot = 0;
    unsigned int hbad = 0;
    VR_CUDA(cudaMemcpyAsync(&htot, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(&hbad, nbad.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    VR_REQUIRE(hbad == 0, VRGPU_EINVAL, "pfof holds ids outside [0, ngroups]");
    gs.ntot = htot;
    if (htot == 0) return;
    DBuf<uint64_t> keys(ctx, htot), keys_tmp(ctx, htot);
    DBuf<uint32_t> vals_tmp(ctx, htot);
    gs.idx.alloc(ctx, htot);
    gs.gid.alloc(ctx, htot);
    VR_LAUNCH(ctx, k_compact_grouped, grid_for(n, T), T, 0, pfof_dev, flags.p, excl.p, n, keys.p, gs.idx.p);
    radix_sort_pairs(ctx, keys.p, gs.idx.p, keys_tmp.p, vals_tmp.p, htot, bits_for_u64((uint64_t)ngroups));
    VR_LAUNCH(ctx, k_keys_to_gid, grid_for(htot, T), T, 0, keys.p, (size_t)htot, gs.gid.p);
}

void gather_groupset(vrgpu_ctx *ctx, const DeviceParticles &P, GroupSet &gs, bool want_vel, bool want_mass)
{
    if (gs.ntot == 0) return;
    const int T = 256;
    gs.x.alloc(ctx, gs.ntot); gs.y.alloc(ctx, gs.ntot); gs.z.alloc(ctx, gs.ntot);
    if (want_vel) {
        VR_REQUIRE(P.has_vel, VRGPU_EINVAL, "velocities required");
        gs.vx.alloc(ctx, gs.ntot); gs.vy.alloc(ctx, gs.ntot); gs.vz.alloc(ctx, gs.ntot);
    }
    if (want_mass) gs.m.alloc(ctx, gs.ntot);
    VR_LAUNCH(ctx, k_gather_cols, grid_for(gs.ntot, T), T, 0, P.pos.p, want_vel ? P.vel.p : (const double *)nullptr,
              P.has_mass ? P.mass.p : (const double *)nullptr, gs.idx.p, gs.ntot, gs.x.p, gs.y.p, gs.z.p,
              want_vel ? gs.vx.p : (double *)nullptr, want_vel ? gs.vy.p : (double *)nullptr, want_vel ? gs.vz.p : (double *)nullptr,
              want_mass ? gs.m.p : (double *)nullptr);
}

// ---------------------------------------------------------------------------------------------------
// Group sizes and member lists for the host (BuildNumInGroup + BuildPGList, src/buildandsortarrays.cxx:11-16, 26-43):
// the group-major slot order IS the concatenation of pglist[1], pglist[2], ... (members in ascending particle index).
// ---------------------------------------------------------------------------------------------------
__global__ void k_pglist_sizes(const uint32_t *__restrict__ off, uint32_t ngroups, long long *__restrict__ numingroup)
{
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g <= ngroups) numingroup[g] = g == 0 ? 0ll : (long long)(off[g + 1] - off[g]);
}
__global__ void k_pglist_index(const uint32_t *__restrict__ idx, size_t ntot, long long *__restrict__ index)
{
    const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s < ntot) index[s] = (long long)idx[s];
}

long long pglist_device(vrgpu_ctx *ctx, const long long *pfof_dev, size_t n, long long ngroups, long long *numingroup_out,
                        long long *index_out)
{
    const int T = 256;
    if (ngroups <= 0 || n == 0) {
        if (numingroup_out) numingroup_out[0] = 0;
        return 0;
    }
    GroupSet gs;
    { KScope k(ctx, "pglist.regroup"); build_groupset(ctx, pfof_dev, n, ngroups, gs); }
    const size_t G1 = (size_t)ngroups + 1;
    DBuf<long long> num(ctx, G1), index(ctx, gs.ntot ? gs.ntot : 1);
    {
        KScope k(ctx, "pglist.widen");
        VR_LAUNCH(ctx, k_pglist_sizes, grid_for(G1, T), T, 0, gs.off.p, (uint32_t)ngroups, num.p);
        if (gs.ntot) VR_LAUNCH(ctx, k_pglist_index, grid_for(gs.ntot, T), T, 0, gs.idx.p, gs.ntot, index.p);
    }
    StageTimer t(ctx);
    if (numingroup_out) download_i64(ctx, num.p, numingroup_out, G1);
    if (index_out && gs.ntot) download_i64(ctx, index.p, index_out, gs.ntot);
    ctx->stats.d2h_ms += t.stop();
    return (long long)gs.ntot;
}

} // namespace vrgpu
