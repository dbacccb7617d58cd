// velociraptor-stf_b200/csrc/energy.cu -- binding energies of finished groups.
//
// Replaces GetBindingEnergy (src/substructureproperties.cxx:3884-4255) for the default reference frame
// (Options::ParticleTypeForRefenceFrame = -1, no per-type mass weights): the caller hands over particles already in group
// order (Part sorted by pfof, numingroup / noffset) together with the centre-of-mass velocities GetCM left in
// pdata[].gcmvel.  Potentials come from the kernels unbinding uses (direct sum below ompunbindnum, kd-tree monopoles from
// there on, like :3922-3940 and :4060-4073), the kinetic energies and every per-group figure (T, Pot, Efrac, the members
// of minimum potential and of minimum total energy) from one pass over the slots with a fixed-shape reduction per group.
//
// Reference quirk kept on purpose: pdata.Pot is the PAIR sum for small groups (:3934, half of the sum of the members'
// potentials) but the plain sum of the members' potentials for large ones (:4186-4196).
#include "groups.cuh"
#include "stages.cuh"

#include <algorithm>
#include <vector>

namespace vrgpu {

constexpr uint32_t EN_BIG = 32768; // members from which a whole CTA reduces one group

struct EnArgs {
    const double *vx, *vy, *vz, *m, *phi;
    const uint32_t *off;     // GroupSet::off, group g (1-based) = slots [off[g], off[g+1])
    const double *gcmvel;    // 3 per group, 0-based
    double *E;               // per slot
    double *T, *Pot, *Efrac; // per group, 0-based
    long long *iminpot, *ibound;
    uint32_t large_threshold;
};

// running figures of one group; `combine` is commutative, so the butterfly leaves the same value in every lane
struct EnAcc {
    double T, P;
    uint32_t nb;
    double pmin, emin;
    uint32_t ipmin, iemin;
};

__device__ __forceinline__ void en_init(EnAcc &a)
{
    a.T = 0.0; a.P = 0.0; a.nb = 0;
    a.pmin = a.emin = __longlong_as_double(0x7ff0000000000000ll);
    a.ipmin = a.iemin = 0xffffffffu;
}
// first strict minimum of a left-to-right scan == smallest index among the minimal values
__device__ __forceinline__ void en_min(double &v, uint32_t &i, double w, uint32_t j)
{
    if (w < v || (w == v && j < i)) { v = w; i = j; }
}
__device__ __forceinline__ void en_member(const EnArgs &A, uint32_t s, uint32_t j, double cx, double cy, double cz, EnAcc &a)
{
    const double m = A.m[s], phi = A.phi[s];
    const double dx = A.vx[s] - cx, dy = A.vy[s] - cy, dz = A.vz[s] - cz;
    double v2 = __dmul_rn(dx, dx);
    v2 = __dadd_rn(v2, __dmul_rn(dy, dy));
    v2 = __dadd_rn(v2, __dmul_rn(dz, dz));
    const double Ti = __dmul_rn(__dmul_rn(0.5, m), v2); // :4040-4049 / :4181-4195
    const double E = __dadd_rn(Ti, phi);
    A.E[s] = E;
    a.T += Ti;
    a.P += phi;
    a.nb += E < 0.0 ? 1u : 0u;
    en_min(a.pmin, a.ipmin, phi, j);
    en_min(a.emin, a.iemin, E, j);
}
__device__ __forceinline__ void en_combine(EnAcc &a, const EnAcc &b)
{
    a.T += b.T;
    a.P += b.P;
    a.nb += b.nb;
    en_min(a.pmin, a.ipmin, b.pmin, b.ipmin);
    en_min(a.emin, a.iemin, b.emin, b.iemin);
}
__device__ __forceinline__ void en_warp_reduce(EnAcc &a)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        EnAcc b;
        b.T = __shfl_xor_sync(0xffffffffu, a.T, o);
        b.P = __shfl_xor_sync(0xffffffffu, a.P, o);
        b.nb = __shfl_xor_sync(0xffffffffu, a.nb, o);
        b.pmin = __shfl_xor_sync(0xffffffffu, a.pmin, o);
        b.emin = __shfl_xor_sync(0xffffffffu, a.emin, o);
        b.ipmin = __shfl_xor_sync(0xffffffffu, a.ipmin, o);
        b.iemin = __shfl_xor_sync(0xffffffffu, a.iemin, o);
        en_combine(a, b);
    }
}
__device__ __forceinline__ void en_store(const EnArgs &A, uint32_t g, uint32_t n, const EnAcc &a)
{
    const bool large = n >= A.large_threshold;
    A.T[g - 1] = a.T;
    A.Pot[g - 1] = large ? a.P : 0.5 * a.P;
    A.Efrac[g - 1] = (double)a.nb / (double)n;
    A.iminpot[g - 1] = (long long)a.ipmin;
    A.ibound[g - 1] = (long long)a.iemin;
}

// one warp per group below EN_BIG members
__global__ void __launch_bounds__(256) k_group_energy(EnArgs A, uint32_t ngroups)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t g = 1 + wg; g <= ngroups; g += nw) {
        const uint32_t s0 = A.off[g], s1 = A.off[g + 1], n = s1 - s0;
        if (n >= EN_BIG) continue; // k_group_energy_big
        if (n == 0) {
            if (lane == 0) { A.T[g - 1] = A.Pot[g - 1] = A.Efrac[g - 1] = 0.0; A.iminpot[g - 1] = A.ibound[g - 1] = 0; }
            continue;
        }
        const double cx = A.gcmvel[3 * (g - 1)], cy = A.gcmvel[3 * (g - 1) + 1], cz = A.gcmvel[3 * (g - 1) + 2];
        EnAcc a;
        en_init(a);
        for (uint32_t s = s0 + lane; s < s1; s += 32) en_member(A, s, s - s0, cx, cy, cz, a);
        en_warp_reduce(a);
        if (lane == 0) en_store(A, g, n, a);
    }
}

// one CTA per group of EN_BIG members or more
__global__ void __launch_bounds__(1024) k_group_energy_big(EnArgs A)
{
    __shared__ EnAcc red[32];
    const uint32_t g = blockIdx.x + 1;
    const uint32_t s0 = A.off[g], s1 = A.off[g + 1], n = s1 - s0;
    if (n < EN_BIG) return;
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const double cx = A.gcmvel[3 * (g - 1)], cy = A.gcmvel[3 * (g - 1) + 1], cz = A.gcmvel[3 * (g - 1) + 2];
    EnAcc a;
    en_init(a);
    for (uint32_t s = s0 + threadIdx.x; s < s1; s += 1024) en_member(A, s, s - s0, cx, cy, cz, a);
    en_warp_reduce(a);
    if (lane == 0) red[wid] = a;
    __syncthreads();
    if (wid == 0) {
        a = red[lane];
        en_warp_reduce(a);
        if (lane == 0) en_store(A, g, n, a);
    }
}

__global__ void k_en_identity(uint32_t *__restrict__ idx, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) idx[i] = i;
}
__global__ void k_en_gid(const uint32_t *__restrict__ off, uint32_t ngroups, uint32_t *__restrict__ gid)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t g = 1 + wg; g <= ngroups; g += nw)
        for (uint32_t s = off[g] + lane; s < off[g + 1]; s += 32) gid[s] = g;
}

void binding_energy_host(vrgpu_ctx *ctx, const DeviceParticles &P, long long ngroups, const long long *offsets_host,
                         const double *gcmvel_host, double G, double eps, double theta, int bucket, long long large_threshold,
                         double *phi_out, double *E_out, double *T_out, double *Pot_out, double *Efrac_out, long long *iminpot_out,
                         long long *ibound_out)
{
    const long long ntot = offsets_host[ngroups];
    VR_REQUIRE(offsets_host[0] == 0, VRGPU_EINVAL, "offsets[0] must be 0");
    VR_REQUIRE(ntot >= 0 && (size_t)ntot <= P.n, VRGPU_EINVAL, "offsets run past the particle view");
    VR_REQUIRE(ntot < 0xfffffff0LL && ngroups < 0x7fffffffLL, VRGPU_EINVAL, "too many grouped particles");
    VR_REQUIRE(large_threshold >= 1 && large_threshold < 0x7fffffffLL, VRGPU_EINVAL, "large_threshold must be >= 1");
    VR_REQUIRE(P.has_vel && P.has_mass, VRGPU_EINVAL, "binding energies need velocities and masses");
    if (ntot == 0) {
        for (long long g = 0; g < ngroups; g++) { T_out[g] = Pot_out[g] = Efrac_out[g] = 0.0; iminpot_out[g] = ibound_out[g] = 0; }
        return;
    }
    const int T = 256;
    const uint32_t NG = (uint32_t)ngroups;
    std::vector<uint32_t> off((size_t)ngroups + 2);
    off[0] = 0;
    for (long long g = 0; g <= ngroups; g++) {
        VR_REQUIRE(g == 0 || offsets_host[g] >= offsets_host[g - 1], VRGPU_EINVAL, "offsets must be non-decreasing");
        off[(size_t)g + 1] = (uint32_t)offsets_host[g];
    }
    GroupSet gs;
    gs.ntot = (size_t)ntot;
    gs.ngroups = ngroups;
    gs.off.alloc(ctx, off.size());
    gs.idx.alloc(ctx, (size_t)ntot);
    gs.gid.alloc(ctx, (size_t)ntot);
    DBuf<double> gcm(ctx, 3 * (size_t)ngroups);
    VR_CUDA(cudaMemcpyAsync(gs.off.p, off.data(), off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(gcm.p, gcmvel_host, 3 * (size_t)ngroups * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    ctx->stats.h2d_bytes += (long long)(off.size() * sizeof(uint32_t) + 3 * (size_t)ngroups * sizeof(double));
    const unsigned warp_grid = std::max(1u, std::min<unsigned>(grid_for((size_t)NG * 32, T), (unsigned)ctx->sm_count * 16u));
    {
        KScope k(ctx, "energy.layout");
        VR_LAUNCH(ctx, k_en_identity, grid_for((size_t)ntot, T), T, 0, gs.idx.p, (uint32_t)ntot);
        VR_LAUNCH(ctx, k_en_gid, warp_grid, T, 0, gs.off.p, NG, gs.gid.p);
        gather_groupset(ctx, P, gs, true, true);
    }
    DBuf<double> phi(ctx, (size_t)ntot);
    {
        StageTimer t(ctx);
        // n < large_threshold: pair loop (:3922-3940); otherwise Potential() on the group (:4060-4073)
        const long long pp = large_threshold - 1;
        { KScope k(ctx, "potential.direct"); ctx->stats.n_pairs_potential += (long long)potential_direct(ctx, gs, G, eps, pp, phi.p); }
        { KScope k(ctx, "potential.tree"); potential_tree(ctx, gs, G, eps, theta, bucket, pp, phi.p); }
        ctx->stats.potential_ms += t.stop();
    }
    DBuf<double> E(ctx, (size_t)ntot), gT(ctx, NG), gPot(ctx, NG), gEfrac(ctx, NG);
    DBuf<long long> gimin(ctx, NG), gibound(ctx, NG);
    EnArgs A{};
    A.vx = gs.vx.p; A.vy = gs.vy.p; A.vz = gs.vz.p; A.m = gs.m.p; A.phi = phi.p;
    A.off = gs.off.p; A.gcmvel = gcm.p; A.E = E.p;
    A.T = gT.p; A.Pot = gPot.p; A.Efrac = gEfrac.p; A.iminpot = gimin.p; A.ibound = gibound.p;
    A.large_threshold = (uint32_t)large_threshold;
    bool any_big = false;
    for (long long g = 0; g < ngroups && !any_big; g++) any_big = offsets_host[g + 1] - offsets_host[g] >= (long long)EN_BIG;
    {
        KScope k(ctx, "energy.groups");
        if (any_big) VR_LAUNCH(ctx, k_group_energy_big, NG, 1024, 0, A);
        VR_LAUNCH(ctx, k_group_energy, warp_grid, T, 0, A, NG);
    }
    {
        StageTimer t(ctx);
        KScope k(ctx, "io.download");
        auto d2h = [&](void *dst, const void *src, size_t bytes) {
            if (!dst || !bytes) return;
            VR_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
            ctx->stats.d2h_bytes += (long long)bytes;
        };
        d2h(phi_out, phi.p, (size_t)ntot * sizeof(double));
        d2h(E_out, E.p, (size_t)ntot * sizeof(double));
        d2h(T_out, gT.p, NG * sizeof(double));
        d2h(Pot_out, gPot.p, NG * sizeof(double));
        d2h(Efrac_out, gEfrac.p, NG * sizeof(double));
        d2h(iminpot_out, gimin.p, NG * sizeof(long long));
        d2h(ibound_out, gibound.p, NG * sizeof(long long));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->stats.d2h_ms += t.stop();
    }
}

// ---------------------------------------------------------------------------------------------------
// Per-group ordering by a key (SortAccordingtoBindingEnergy's qsort by DenCompare / PotCompare,
// substructureproperties.cxx:4302-4310): two stable LSD radix sorts -- by the order-preserving image of the key, then by
// group id -- leave every group's members in ascending key order, equal keys in ascending member order.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t en_ordered_bits(double d)
{
    const uint64_t u = (uint64_t)__double_as_longlong(d);
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}
__global__ void k_en_sortkeys(const double *__restrict__ key, uint32_t ntot, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot) return;
    keys[s] = en_ordered_bits(__dadd_rn(key[s], 0.0)); // -0.0 + 0.0 = +0.0: the two zeros compare equal, like in DenCompare
    vals[s] = s;
}
__global__ void k_en_gidkeys(const uint32_t *__restrict__ gid, const uint32_t *__restrict__ vals, uint32_t ntot, uint64_t *__restrict__ keys)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < ntot) keys[p] = gid[vals[p]];
}
// position p of the sorted list lies in the same group as slot p (groups are contiguous and keep their sizes)
__global__ void k_en_order_out(const uint32_t *__restrict__ vals, const uint32_t *__restrict__ gid, const uint32_t *__restrict__ off,
                               uint32_t ntot, long long *__restrict__ order)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < ntot) order[p] = (long long)(vals[p] - off[gid[p]]);
}

void sort_groups_host(vrgpu_ctx *ctx, long long ngroups, const long long *offsets_host, const double *key_host, long long *order_out)
{
    const long long ntot = offsets_host[ngroups];
    VR_REQUIRE(offsets_host[0] == 0, VRGPU_EINVAL, "offsets[0] must be 0");
    VR_REQUIRE(ntot >= 0 && ntot < 0xfffffff0LL && ngroups < 0x7fffffffLL, VRGPU_EINVAL, "too many grouped particles");
    if (ntot == 0) return;
    const int T = 256;
    const uint32_t NG = (uint32_t)ngroups, N = (uint32_t)ntot;
    std::vector<uint32_t> off((size_t)ngroups + 2);
    off[0] = 0;
    for (long long g = 0; g <= ngroups; g++) {
        VR_REQUIRE(g == 0 || offsets_host[g] >= offsets_host[g - 1], VRGPU_EINVAL, "offsets must be non-decreasing");
        off[(size_t)g + 1] = (uint32_t)offsets_host[g];
    }
    DBuf<uint32_t> doff(ctx, off.size()), gid(ctx, N), vals(ctx, N), vals_tmp(ctx, N);
    DBuf<uint64_t> keys(ctx, N), keys_tmp(ctx, N);
    DBuf<double> key(ctx, N);
    DBuf<long long> order(ctx, N);
    {
        StageTimer t(ctx);
        KScope k(ctx, "io.upload");
        VR_CUDA(cudaMemcpyAsync(doff.p, off.data(), off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(key.p, key_host, (size_t)N * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        ctx->stats.h2d_bytes += (long long)(off.size() * sizeof(uint32_t) + (size_t)N * sizeof(double));
        ctx->stats.h2d_ms += t.stop();
    }
    {
        KScope k(ctx, "energy.sort");
        const unsigned warp_grid = std::max(1u, std::min<unsigned>(grid_for((size_t)NG * 32, T), (unsigned)ctx->sm_count * 16u));
        VR_LAUNCH(ctx, k_en_gid, warp_grid, T, 0, doff.p, NG, gid.p);
        VR_LAUNCH(ctx, k_en_sortkeys, grid_for(N, T), T, 0, key.p, N, keys.p, vals.p);
        radix_sort_pairs(ctx, keys.p, vals.p, keys_tmp.p, vals_tmp.p, N, 64);
        VR_LAUNCH(ctx, k_en_gidkeys, grid_for(N, T), T, 0, gid.p, vals.p, N, keys.p);
        int gbits = 1;
        while ((1ull << gbits) <= (unsigned long long)NG) gbits++;
        radix_sort_pairs(ctx, keys.p, vals.p, keys_tmp.p, vals_tmp.p, N, gbits);
        VR_LAUNCH(ctx, k_en_order_out, grid_for(N, T), T, 0, vals.p, gid.p, doff.p, N, order.p);
    }
    {
        StageTimer t(ctx);
        download_i64(ctx, order.p, order_out, N);
        ctx->stats.d2h_ms += t.stop();
    }
}

} // namespace vrgpu
