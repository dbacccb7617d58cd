// velociraptor-stf_b200/csrc/unbind.cu -- CheckUnboundGroups / Unbind on the device.
//
// Replaces src/unbind.cxx:196-304 (CheckUnboundGroups), :320-729 (Unbind) and their helpers
// GetBoundFractionAndMaxE (:146-185), CheckGroupForBoundness (:49-60), FillUnboundArrays (:62-98),
// UpdateCMForUnboundParticles (:111-127), RemoveUnboundParticles (:101-108), RemoveGroup (:129-140)
// for the configuration every sample cfg uses: Frac_pot_ref = 1 (CM frame of the whole group) and
// Keep_background_potential = 1 (potentials computed once).
//
// Structure (mirrors the reference's two loops):
//   phase A  groups with n >= large_threshold (ompunbindnum): the reference's "large" loop can only ever make
//            ONE removal pass (its continue-condition contradicts FillUnboundArrays' cancel rule), so this is a
//            data-parallel pass: energies -> per-group counters -> global (gid, E) sort -> decide -> apply.
//   phase B  every group that now has 0 < n <= large_threshold: one CTA per group runs the reference's
//            small-group state machine entirely in shared memory (energies, bitonic sort by E, tail removal,
//            CM update, termination rules).
// Deviation (documented in DESIGN.md): `oldnumingroup`, which the reference reads uninitialised in
// the small-group loop, is initialised to the group's size when its loop starts.
#include "groups.cuh"
#include "stages.cuh"

#include <algorithm>
#include <cmath>

namespace vrgpu {

constexpr int UB_CAP = 2048; // max members handled by the in-CTA state machine
constexpr int UB_T = 256;

struct UbParams {
    double Eratio, minEfrac, maxunbindfrac, maxunboundfrac, maxallowed, marginal_tol;
    int unbindtype;
    long long min_size, large_threshold;
    int fixed_frame; // Frac_pot_ref != 1: the velocity frame does not follow the removals (unbind.cxx:118)
};

struct UbArrays {
    const double *vx, *vy, *vz, *m, *phi;
    const uint32_t *gid, *off;
    double *E;                // per slot
    uint32_t *ord;            // per slot: arrangement of each group's segment (gPart order)
    unsigned char *alive;     // per slot
    unsigned char *marg;      // per slot: |E| <= tol |phi| at the last evaluation
    double *cmvel;            // 3 per group
    double *gmass;            // per group
    uint32_t *ning;           // per group: current member count
    uint32_t *nunb, *nefrac;  // per group counters (phase A)
    unsigned long long *maxE; // per group, ordered-uint image of a double (phase A)
    uint32_t *nrem;           // per group (phase A)
    int *status;              // per group: 0 untouched, 1 removed some, 2 dissolved
    unsigned int *changed;    // global flag (iunbindflag != 0)
};

__device__ __forceinline__ unsigned long long dbl_to_ord(double d)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(d);
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double ord_to_dbl(unsigned long long u)
{
    u = (u & 0x8000000000000000ull) ? (u & 0x7fffffffffffffffull) : ~u;
    return __longlong_as_double((long long)u);
}

// energy of one member in the frame cm (GetBoundFractionAndMaxE, unbind.cxx:163-178)
__device__ __forceinline__ void member_energy(const UbArrays &A, const UbParams &P, uint32_t slot, double cx, double cy, double cz,
                                              double &E, bool &bound_frac, bool &marginal)
{
    double m = A.m[slot];
    double dx = A.vx[slot] - cx, dy = A.vy[slot] - cy, dz = A.vz[slot] - cz;
    double v2 = 0.0;
    v2 = __dadd_rn(v2, __dmul_rn(dx, dx));
    v2 = __dadd_rn(v2, __dmul_rn(dy, dy));
    v2 = __dadd_rn(v2, __dmul_rn(dz, dz));
    double Ti = __dmul_rn(__dmul_rn(0.5, m), v2);
    double phi = A.phi[slot];
    E = __dadd_rn(__dmul_rn(P.Eratio, Ti), phi);
    bound_frac = __dadd_rn(Ti, phi) < 0;
    marginal = fabs(E) <= P.marginal_tol * fabs(phi);
}

// CM velocity and mass of every group (unbind.cxx:447-455): one warp per group, fixed reduction shape
constexpr uint32_t CM_BIG = 32768; // members from which a whole CTA sums one group (fixed order: deterministic)

__global__ void __launch_bounds__(1024) k_group_cm_big(UbArrays A)
{
    __shared__ double red[4][32];
    const uint32_t g = blockIdx.x + 1;
    const uint32_t s0 = A.off[g], s1 = A.off[g + 1];
    if (s1 - s0 < CM_BIG) return;
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    double sm = 0, sx = 0, sy = 0, sz = 0;
    for (uint32_t s = s0 + threadIdx.x; s < s1; s += 1024) {
        const double m = A.m[s];
        sm += m; sx += A.vx[s] * m; sy += A.vy[s] * m; sz += A.vz[s] * m;
    }
    for (int o = 16; o; o >>= 1) {
        sm += __shfl_xor_sync(0xffffffffu, sm, o); sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o); sz += __shfl_xor_sync(0xffffffffu, sz, o);
    }
    if (lane == 0) { red[0][wid] = sm; red[1][wid] = sx; red[2][wid] = sy; red[3][wid] = sz; }
    __syncthreads();
    if (wid == 0) {
        sm = red[0][lane]; sx = red[1][lane]; sy = red[2][lane]; sz = red[3][lane];
        for (int o = 16; o; o >>= 1) {
            sm += __shfl_xor_sync(0xffffffffu, sm, o); sx += __shfl_xor_sync(0xffffffffu, sx, o);
            sy += __shfl_xor_sync(0xffffffffu, sy, o); sz += __shfl_xor_sync(0xffffffffu, sz, o);
        }
        if (lane == 0) {
            const double inv = 1.0 / sm;
            A.gmass[g] = sm;
            A.cmvel[3 * g] = sx * inv; A.cmvel[3 * g + 1] = sy * inv; A.cmvel[3 * g + 2] = sz * inv;
            A.ning[g] = s1 - s0;
            A.status[g] = 0;
        }
    }
}

__global__ void __launch_bounds__(256) k_group_cm(UbArrays A, uint32_t ngroups)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t g = 1 + wg; g <= ngroups; g += nw) {
        uint32_t s0 = A.off[g], s1 = A.off[g + 1];
        if (s1 - s0 >= CM_BIG) continue; // k_group_cm_big
        double sm = 0, sx = 0, sy = 0, sz = 0;
        for (uint32_t s = s0 + lane; s < s1; s += 32) {
            double m = A.m[s];
            sm += m; sx += A.vx[s] * m; sy += A.vy[s] * m; sz += A.vz[s] * m;
        }
        for (int o = 16; o; o >>= 1) {
            sm += __shfl_xor_sync(0xffffffffu, sm, o); sx += __shfl_xor_sync(0xffffffffu, sx, o);
            sy += __shfl_xor_sync(0xffffffffu, sy, o); sz += __shfl_xor_sync(0xffffffffu, sz, o);
        }
        if (lane == 0) {
            double inv = 1.0 / sm;
            A.gmass[g] = sm;
            A.cmvel[3 * g] = sx * inv; A.cmvel[3 * g + 1] = sy * inv; A.cmvel[3 * g + 2] = sz * inv;
            A.ning[g] = s1 - s0;
            A.status[g] = 0;
        }
    }
}

__global__ void k_init_slots(UbArrays A, uint32_t ntot)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot) return;
    A.ord[s] = s;
    A.alive[s] = 1;
    A.marg[s] = 0;
}

// ---------------------------------------------------------------------------------------------------
// phase A: large groups
// ---------------------------------------------------------------------------------------------------
__global__ void k_large_energy(UbArrays A, UbParams P, uint32_t ntot, uint32_t *__restrict__ lflag)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t g = 0;
    bool large = false;
    if (s < ntot) {
        g = A.gid[s];
        large = (long long)(A.off[g + 1] - A.off[g]) >= P.large_threshold;
        lflag[s] = large ? 1u : 0u;
    }
    double E = 0;
    bool bf = false, mg = false;
    if (large) {
        member_energy(A, P, s, A.cmvel[3 * g], A.cmvel[3 * g + 1], A.cmvel[3 * g + 2], E, bf, mg);
        A.E[s] = E;
        A.marg[s] = mg ? 1 : 0;
    }
    // warp-aggregated per-group counters (integer sums and a max: order independent)
    const unsigned key = large ? g : 0u;
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const bool uni = __all_sync(0xffffffffu, peers == 0xffffffffu);
    if (uni) {
        if (large) { // whole warp inside one large group
            const unsigned lane = threadIdx.x & 31;
            unsigned u = __ballot_sync(0xffffffffu, E > 0), f = __ballot_sync(0xffffffffu, bf);
            unsigned long long mo = dbl_to_ord(E);
            for (int o = 16; o; o >>= 1) {
                unsigned long long y = __shfl_xor_sync(0xffffffffu, mo, o);
                if (y > mo) mo = y;
            }
            if (lane == 0) { atomicAdd(&A.nunb[g], __popc(u)); atomicAdd(&A.nefrac[g], __popc(f)); atomicMax(&A.maxE[g], mo); }
        }
    } else if (large) {
        // mixed warp (group boundary or array tail): per-lane atomics
        atomicAdd(&A.nunb[g], E > 0 ? 1u : 0u);
        atomicAdd(&A.nefrac[g], bf ? 1u : 0u);
        atomicMax(&A.maxE[g], dbl_to_ord(E));
    }
}

__global__ void k_large_keys(const UbArrays A, const uint32_t *__restrict__ lflag, const uint32_t *__restrict__ lexcl, uint32_t ntot,
                             uint64_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot || !lflag[s]) return;
    uint32_t k = lexcl[s];
    keys[k] = dbl_to_ord(A.E[s]);
    vals[k] = s;
}
__global__ void k_large_gidkeys(const UbArrays A, const uint32_t *__restrict__ vals, uint32_t nl, uint64_t *__restrict__ keys)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < nl) keys[k] = A.gid[vals[k]];
}
__global__ void k_large_lstart(const uint32_t *__restrict__ off, const uint32_t *__restrict__ lexcl, uint32_t ngroups, uint32_t ntot,
                               uint32_t *__restrict__ lstart)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g > ngroups) return;
    uint32_t s = off[g];
    lstart[g] = s < ntot ? lexcl[s] : 0u;
}

// one thread per group: the decisions of unbind.cxx:556-570 (first evaluation of a large group)
__global__ void k_large_decide(UbArrays A, UbParams P, uint32_t ngroups)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (g > ngroups) return;
    uint32_t n = A.off[g + 1] - A.off[g];
    A.nrem[g] = 0;
    if ((long long)n < P.large_threshold) return;
    uint32_t nunb = A.nunb[g];
    double Efrac = (double)A.nefrac[g] / (double)n;
    double maxE = ord_to_dbl(A.maxE[g]);
    if ((double)nunb >= P.maxunboundfrac * (double)n) {
        A.status[g] = 2;
        A.nrem[g] = n;
        atomicOr(A.changed, 1u);
        return;
    }
    long long maxunbindsize = (long long)(P.maxunbindfrac * (double)nunb + 1);
    bool check;
    if (P.unbindtype == 0) check = ((Efrac < P.minEfrac) || (maxE > 0)) && ((long long)n >= P.min_size);
    else check = (maxE > 0) && ((long long)n >= P.min_size);
    long long nEplus = 0;
    if (check) {
        // FillUnboundArrays on an energy-sorted group: the tail walk takes max(nunbound, nEfrac) entries, capped
        long long nEfrac = 0;
        if (P.unbindtype == 0 && Efrac < P.minEfrac) nEfrac = (long long)((P.minEfrac - Efrac) * (double)n);
        long long want = P.unbindtype == 1 ? (long long)nunb : ((long long)nunb > nEfrac ? (long long)nunb : nEfrac);
        nEplus = want < maxunbindsize ? want : maxunbindsize;
        if (nEplus > (long long)n) nEplus = n;
        if ((double)nEplus < P.maxallowed * (double)n) nEplus = 0;
    }
    if (nEplus > 0) {
        A.status[g] = 1;
        A.nrem[g] = (uint32_t)nEplus;
        atomicOr(A.changed, 1u);
    }
}

// one thread per sorted large-group slot: record the arrangement and retire the removed tail
__global__ void k_large_apply(UbArrays A, const uint32_t *__restrict__ sorted_slots, const uint32_t *__restrict__ lstart, uint32_t nl)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nl) return;
    uint32_t s = sorted_slots[k];
    uint32_t g = A.gid[s];
    uint32_t p = k - lstart[g];
    uint32_t n = A.off[g + 1] - A.off[g];
    A.ord[A.off[g] + p] = s;
    if (A.status[g] == 2 || p >= n - A.nrem[g]) A.alive[s] = 0;
}

// one warp per large group that lost particles: UpdateCMForUnboundParticles (unbind.cxx:111-127), RemoveGroup (:129-140)
__global__ void __launch_bounds__(256) k_large_finish(UbArrays A, UbParams P, uint32_t ngroups)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t g = 1 + wg; g <= ngroups; g += nw) {
        int st = A.status[g];
        if (st == 0) continue;
        uint32_t n = A.off[g + 1] - A.off[g];
        if ((long long)n < P.large_threshold) continue;
        uint32_t base = A.off[g];
        if (st == 2) { if (lane == 0) A.ning[g] = 0; continue; }
        uint32_t nrem = A.nrem[g], nn = n - nrem;
        double sm = 0, sx = 0, sy = 0, sz = 0;
        for (uint32_t p = nn + lane; p < n; p += 32) {
            uint32_t s = A.ord[base + p];
            double m = A.m[s];
            sm += m; sx += A.vx[s] * m; sy += A.vy[s] * m; sz += A.vz[s] * m;
        }
        for (int o = 16; o; o >>= 1) {
            sm += __shfl_xor_sync(0xffffffffu, sm, o); sx += __shfl_xor_sync(0xffffffffu, sx, o);
            sy += __shfl_xor_sync(0xffffffffu, sy, o); sz += __shfl_xor_sync(0xffffffffu, sz, o);
        }
        if ((long long)nn < P.min_size) {
            for (uint32_t p = lane; p < nn; p += 32) A.alive[A.ord[base + p]] = 0;
            nn = 0;
        }
        if (lane == 0) {
            if (!P.fixed_frame) {
                double mass = A.gmass[g];
                double temp = 1.0 / mass;
                double cx = A.cmvel[3 * g] - sx * temp, cy = A.cmvel[3 * g + 1] - sy * temp, cz = A.cmvel[3 * g + 2] - sz * temp;
                temp = mass / (mass - sm);
                A.cmvel[3 * g] = cx * temp; A.cmvel[3 * g + 1] = cy * temp; A.cmvel[3 * g + 2] = cz * temp;
                A.gmass[g] = mass - sm;
            }
            A.ning[g] = nn;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// phase B: one CTA per group with 0 < n <= large_threshold -- the small-group loop of unbind.cxx:641-720
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool ub_less(double ea, uint32_t ia, double eb, uint32_t ib)
{
    return ea < eb || (ea == eb && ia < ib);
}

template <typename T> __device__ __forceinline__ T ub_block_sum(T v, T *scratch)
{
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    T r = 0;
    for (int w = 0; w < UB_T / 32; w++) r += scratch[w];
    return r;
}
__device__ __forceinline__ double ub_block_max(double v, double *scratch)
{
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    double r = scratch[0];
    for (int w = 1; w < UB_T / 32; w++) r = fmax(r, scratch[w]);
    return r;
}

// Keep_background_potential = 0 (unbind.cxx:672-690): before the unbound tail leaves, every member loses the pair terms it
// shares with the removed particles.  The terms are taken off one removed particle after the other, in the order of
// nEplusid[] (from the tail of the arrangement towards its head), with the reference's operation order, so the updated
// potentials -- and with them every later energy comparison -- are the reference's bit for bit.
struct UbBg {
    const double *x, *y, *z; // group-major positions
    double *phi;             // the potentials k_unbind_small reads through A.phi (same buffer)
    double G, eps2;
};

template <bool BG0> __global__ void __launch_bounds__(UB_T) k_unbind_small(UbArrays A, UbParams P, uint32_t ngroups, UbBg B)
{
    __shared__ double sE[UB_CAP];
    __shared__ uint32_t sid[UB_CAP];
    __shared__ double dscr[UB_T / 32];
    __shared__ int iscr[UB_T / 32];
    __shared__ long long s_nEplus;
    const uint32_t g = blockIdx.x + 1;
    if (g > ngroups) return;
    uint32_t n = A.ning[g];
    if (n == 0 || (long long)n > P.large_threshold || n > UB_CAP) return;
    const uint32_t base = A.off[g];
    const int tid = threadIdx.x;
    for (uint32_t j = tid; j < n; j += UB_T) sid[j] = A.ord[base + j];
    double cx = A.cmvel[3 * g], cy = A.cmvel[3 * g + 1], cz = A.cmvel[3 * g + 2], mass = A.gmass[g];
    long long oldn = n; // intended initialisation of `oldnumingroup`
    bool changed = false, sortflag = true, first = true, check = true;
    __syncthreads();
    while (check) {
        // ---- GetBoundFractionAndMaxE (unbind.cxx:146-185) ----
        int nunb_l = 0, nef_l = 0;
        double mx_l = -1e32;
        for (uint32_t j = tid; j < n; j += UB_T) {
            double E;
            bool bf, mg;
            member_energy(A, P, sid[j], cx, cy, cz, E, bf, mg);
            sE[j] = E;
            A.marg[sid[j]] = mg ? 1 : 0;
            nunb_l += E > 0;
            nef_l += bf;
            mx_l = fmax(mx_l, E);
        }
        const int nunb = ub_block_sum<int>(nunb_l, iscr);
        const int nef = ub_block_sum<int>(nef_l, iscr);
        const double maxE = ub_block_max(mx_l, dscr);
        const double Efrac = (double)nef / (double)n;
        bool sorted_now = false;
        if ((double)nunb < P.maxunboundfrac * (double)n && sortflag) {
            // std::sort by energy (unbind.cxx:181-184): bitonic network over the next power of two
            uint32_t np2 = 1;
            while (np2 < n) np2 <<= 1;
            for (uint32_t j = n + tid; j < np2; j += UB_T) { sE[j] = 1e308; sid[j] = 0xffffffffu; }
            __syncthreads();
            for (uint32_t k = 2; k <= np2; k <<= 1)
                for (uint32_t jj = k >> 1; jj > 0; jj >>= 1) {
                    for (uint32_t i = tid; i < np2; i += UB_T) {
                        uint32_t l = i ^ jj;
                        if (l > i) {
                            bool up = (i & k) == 0;
                            double ei = sE[i], el = sE[l];
                            uint32_t ii = sid[i], il = sid[l];
                            bool sw = up ? ub_less(el, il, ei, ii) : ub_less(ei, ii, el, il);
                            if (sw) { sE[i] = el; sE[l] = ei; sid[i] = il; sid[l] = ii; }
                        }
                    }
                    __syncthreads();
                }
            sorted_now = true;
        }
        __syncthreads();
        if (first) {
            first = false;
            if ((double)nunb >= P.maxunboundfrac * (double)n) { // unbind.cxx:654-658
                for (uint32_t j = tid; j < n; j += UB_T) A.alive[sid[j]] = 0;
                n = 0;
                changed = true;
                break;
            }
        }
        // ---- CheckGroupForBoundness + FillUnboundArrays (unbind.cxx:49-98) ----
        const long long maxunbindsize = (long long)(P.maxunbindfrac * (double)nunb + 1);
        if (P.unbindtype == 0) check = ((Efrac < P.minEfrac) || (maxE > 0)) && ((long long)n >= P.min_size);
        else check = (maxE > 0) && ((long long)n >= P.min_size);
        long long nEplus = 0;
        if (check) {
            long long nEfrac = 0;
            if (P.unbindtype == 0 && Efrac < P.minEfrac) nEfrac = (long long)((P.minEfrac - Efrac) * (double)n);
            if (sorted_now) {
                long long want = P.unbindtype == 1 ? (long long)nunb : ((long long)nunb > nEfrac ? (long long)nunb : nEfrac);
                nEplus = want < maxunbindsize ? want : maxunbindsize;
                if (nEplus > (long long)n) nEplus = n;
            } else {
                // stale arrangement (sortflag == false): literal walk from the tail
                if (tid == 0) {
                    long long c = 0;
                    for (long long j = 0; j < maxunbindsize && j < (long long)n; j++) {
                        double e = sE[n - 1 - j];
                        if (e > 0 || (P.unbindtype == 0 && c < nEfrac)) c++;
                        else break;
                    }
                    s_nEplus = c;
                }
                __syncthreads();
                nEplus = s_nEplus;
                __syncthreads();
            }
            if ((double)nEplus < P.maxallowed * (double)n) { nEplus = 0; check = false; }
        }
        if (!check) break;
        // ---- loop body (unbind.cxx:666-709) ----
        changed = true;
        {
            double sm_l = 0, sx_l = 0, sy_l = 0, sz_l = 0;
            for (uint32_t p = n - (uint32_t)nEplus + tid; p < n; p += UB_T) {
                uint32_t s = sid[p];
                double m = A.m[s];
                sm_l += m; sx_l += A.vx[s] * m; sy_l += A.vy[s] * m; sz_l += A.vz[s] * m;
                A.alive[s] = 0;
            }
            double sm = ub_block_sum<double>(sm_l, dscr), sx = ub_block_sum<double>(sx_l, dscr);
            double sy = ub_block_sum<double>(sy_l, dscr), sz = ub_block_sum<double>(sz_l, dscr);
            if (!P.fixed_frame) {
                double temp = 1.0 / mass;
                cx -= sx * temp; cy -= sy * temp; cz -= sz * temp;
                temp = mass / (mass - sm);
                cx *= temp; cy *= temp; cz *= temp;
                mass -= sm;
            }
        }
        if (BG0) {
            const uint32_t nrm = (uint32_t)nEplus;
            for (uint32_t j = tid; j < n; j += UB_T) {
                const uint32_t sj = sid[j];
                const double xj = B.x[sj], yj = B.y[sj], zj = B.z[sj], mj = A.m[sj];
                double ph = B.phi[sj];
                for (uint32_t k = 0; k < nrm; k++) {
                    const uint32_t pk = n - 1 - k; // nEplusid[k] (FillUnboundArrays, unbind.cxx:62-98)
                    if (pk == j) continue;
                    const uint32_t sk = sid[pk];
                    const double dx = __dsub_rn(B.x[sk], xj), dy = __dsub_rn(B.y[sk], yj), dz = __dsub_rn(B.z[sk], zj);
                    double r2 = __dmul_rn(dx, dx);
                    r2 = __dadd_rn(r2, __dmul_rn(dy, dy));
                    r2 = __dadd_rn(r2, __dmul_rn(dz, dz));
                    r2 = __dadd_rn(r2, B.eps2);
                    const double rinv = __ddiv_rn(1.0, __dsqrt_rn(r2));
                    ph = __dadd_rn(ph, __dmul_rn(__dmul_rn(B.G, __dmul_rn(A.m[sk], mj)), rinv));
                }
                B.phi[sj] = ph;
            }
            __syncthreads();
        }
        n -= (uint32_t)nEplus;
        if ((double)nEplus < P.maxallowed * (double)n) break; // unbind.cxx:695-697
        sortflag = false;
        if ((double)(oldn - (long long)n) > P.maxallowed * (double)oldn) { oldn = n; sortflag = true; }
        __syncthreads();
    }
    // RemoveGroup (unbind.cxx:129-140)
    if (n > 0 && (long long)n < P.min_size) {
        for (uint32_t j = tid; j < n; j += UB_T) A.alive[sid[j]] = 0;
        n = 0;
        changed = true;
    }
    if (tid == 0) {
        A.ning[g] = n;
        if (changed) atomicOr(A.changed, 1u);
    }
}

// ---------------------------------------------------------------------------------------------------
// relabelling after unbinding (ReorderGroupIDs, buildandsortarrays.cxx:170-181; canonical tie break)
// ---------------------------------------------------------------------------------------------------
__global__ void k_iota_u32(uint32_t *a, uint32_t n)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = i;
}
// slots are group-major, so the lanes of a warp hold runs of equal group id: one segmented reduction per run, then one
// atomic per (warp, group) instead of one per member
__global__ void k_new_group_stats(const UbArrays A, const uint32_t *__restrict__ idx, const long long *__restrict__ pid, uint32_t ntot,
                                  uint32_t *__restrict__ newsize, unsigned long long *__restrict__ minpid, unsigned long long *__restrict__ nmarg)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    bool mg = false;
    uint32_t g = 0xffffffffu, cnt = 0;
    unsigned long long p = ~0ull;
    if (s < ntot) {
        mg = A.marg[s];
        g = A.gid[s];
        if (A.alive[s]) {
            const uint32_t i = idx[s];
            cnt = 1;
            p = pid ? (unsigned long long)pid[i] : (unsigned long long)i;
        }
    }
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t go = __shfl_down_sync(0xffffffffu, g, o);
        const uint32_t co = __shfl_down_sync(0xffffffffu, cnt, o);
        const unsigned long long po = __shfl_down_sync(0xffffffffu, p, o);
        if (lane + o < 32 && go == g) { cnt += co; p = po < p ? po : p; }
    }
    const uint32_t gprev = __shfl_up_sync(0xffffffffu, g, 1);
    if (s < ntot && cnt > 0 && (lane == 0 || gprev != g)) {
        atomicAdd(&newsize[g], cnt);
        atomicMin(&minpid[g], p);
    }
    unsigned b = __ballot_sync(0xffffffffu, mg);
    if (lane == 0 && b) atomicAdd(nmarg, (unsigned long long)__popc(b));
}
__global__ void k_write_unbound(const UbArrays A, const uint32_t *__restrict__ idx, const uint32_t *__restrict__ gid_of, uint32_t ntot,
                                long long *__restrict__ pfof)
{
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot) return;
    pfof[idx[s]] = A.alive[s] ? (long long)gid_of[A.gid[s]] : 0ll;
}
__global__ void k_csr_gid(const uint32_t *__restrict__ off, uint32_t ngroups, uint32_t *__restrict__ gid)
{
    // one warp per group paints its slots
    const unsigned lane = threadIdx.x & 31;
    const uint32_t wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t g = 1 + wg; g <= ngroups; g += nw)
        for (uint32_t s = off[g] + lane; s < off[g + 1]; s += 32) gid[s] = g;
}

static void compute_potentials(vrgpu_ctx *ctx, const GroupSet &gs, double G, double eps, double theta, int bucket, long long pp_threshold,
                               double *phi)
{
    StageTimer t(ctx);
    { KScope k(ctx, "potential.direct"); ctx->stats.n_pairs_potential += (long long)potential_direct(ctx, gs, G, eps, pp_threshold, phi); }
    if (pp_threshold >= 0) { KScope k(ctx, "potential.tree"); potential_tree(ctx, gs, G, eps, theta, bucket, pp_threshold, phi); }
    ctx->stats.potential_ms += t.stop();
}

// ---- kinetic reference frame for Frac_pot_ref != 1 (unbind.cxx:460-544) ---------------------------------------------------
// centre of every group: its centre of mass (CMVELREF) or the position of the member with the lowest potential (POTREF, the
// first strict minimum); one CTA per group, fixed reduction shape
constexpr int FR_T = 128;
__global__ void __launch_bounds__(FR_T) k_frame_centre(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                                                       const double *__restrict__ m, const double *__restrict__ phi, const uint32_t *__restrict__ off,
                                                       int potref, double *__restrict__ centre, double *__restrict__ gmass)
{
    __shared__ double red[4][FR_T];
    __shared__ uint32_t ri[FR_T];
    const uint32_t g = blockIdx.x + 1, s0 = off[g], s1 = off[g + 1];
    const unsigned tid = threadIdx.x;
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    uint32_t best = 0xffffffffu;
    double bphi = 0;
    for (uint32_t s = s0 + tid; s < s1; s += FR_T) {
        const double mm = m[s];
        a0 += x[s] * mm; a1 += y[s] * mm; a2 += z[s] * mm; a3 += mm;
        const double p = phi[s];
        if (best == 0xffffffffu || p < bphi) { best = s; bphi = p; } // slots ascend: the first strict minimum of this thread
    }
    red[0][tid] = potref ? bphi : a0; red[1][tid] = a1; red[2][tid] = a2; red[3][tid] = a3; ri[tid] = best;
    __syncthreads();
    for (int o = FR_T / 2; o; o >>= 1) {
        if ((int)tid < o) {
            if (potref) {
                const uint32_t b = ri[tid + o];
                if (b != 0xffffffffu && (ri[tid] == 0xffffffffu || red[0][tid + o] < red[0][tid] || (red[0][tid + o] == red[0][tid] && b < ri[tid]))) {
                    ri[tid] = b; red[0][tid] = red[0][tid + o];
                }
            } else red[0][tid] += red[0][tid + o];
            red[1][tid] += red[1][tid + o]; red[2][tid] += red[2][tid + o]; red[3][tid] += red[3][tid + o];
        }
        __syncthreads();
    }
    if (tid == 0 && s1 > s0) {
        gmass[g] = red[3][0];
        if (potref) { const uint32_t s = ri[0]; centre[3 * g] = x[s]; centre[3 * g + 1] = y[s]; centre[3 * g + 2] = z[s]; }
        else { const double inv = 1.0 / red[3][0]; centre[3 * g] = red[0][0] * inv; centre[3 * g + 1] = red[1][0] * inv; centre[3 * g + 2] = red[2][0] * inv; }
    }
}
// Particle::Radius() about the group's centre, as a sort key (radii are non-negative: the bit pattern orders them)
__global__ void k_frame_radius(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                               const uint32_t *__restrict__ gid, const double *__restrict__ centre, uint32_t ntot, uint64_t *__restrict__ keys,
                               uint32_t *__restrict__ vals)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ntot) return;
    const uint32_t g = gid[s];
    const double dx = __dsub_rn(x[s], centre[3 * g]), dy = __dsub_rn(y[s], centre[3 * g + 1]), dz = __dsub_rn(z[s], centre[3 * g + 2]);
    const double r = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
    keys[s] = (uint64_t)__double_as_longlong(r);
    vals[s] = s;
}
__global__ void k_frame_gidkeys(const uint32_t *__restrict__ gid, const uint32_t *__restrict__ vals, uint32_t ntot, uint64_t *__restrict__ keys)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < ntot) keys[k] = gid[vals[k]];
}
// mean velocity of the npot innermost members (positions [off[g], off[g] + npot) of the radius order)
__global__ void __launch_bounds__(FR_T) k_frame_cmvel(UbArrays A, const uint32_t *__restrict__ order, double frac, long long npotref)
{
    __shared__ double red[4][FR_T];
    const uint32_t g = blockIdx.x + 1, s0 = A.off[g], n = A.off[g + 1] - s0;
    const unsigned tid = threadIdx.x;
    long long npot = (long long)(frac * (double)n);
    if (npot < npotref) npot = npotref;
    if (npot > (long long)n) npot = n;
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    for (uint32_t q = tid; q < (uint32_t)npot; q += FR_T) {
        const uint32_t s = order[s0 + q];
        const double mm = A.m[s];
        a0 += A.vx[s] * mm; a1 += A.vy[s] * mm; a2 += A.vz[s] * mm; a3 += mm;
    }
    red[0][tid] = a0; red[1][tid] = a1; red[2][tid] = a2; red[3][tid] = a3;
    __syncthreads();
    for (int o = FR_T / 2; o; o >>= 1) {
        if ((int)tid < o) { red[0][tid] += red[0][tid + o]; red[1][tid] += red[1][tid + o]; red[2][tid] += red[2][tid + o]; red[3][tid] += red[3][tid + o]; }
        __syncthreads();
    }
    if (tid == 0 && n > 0) {
        A.cmvel[3 * g] = red[0][0] / red[3][0]; A.cmvel[3 * g + 1] = red[1][0] / red[3][0]; A.cmvel[3 * g + 2] = red[2][0] / red[3][0];
        A.ning[g] = n;
        A.status[g] = 0;
    }
}

void unbind_device(vrgpu_ctx *ctx, const DeviceParticles &P, long long *pfof_dev, long long &ngroups, const vrgpu_unbind_opts &o,
                   int &changed, long long &n_marginal)
{
    changed = 0;
    n_marginal = 0;
    if (ngroups <= 0 || P.n == 0) return;
    VR_REQUIRE(o.fracpotref > 0 && o.fracpotref <= 1.0, VRGPU_EINVAL, "Frac_pot_ref must be in (0, 1]");
    VR_REQUIRE(o.fracpotref == 1.0 || o.cmvelreftype == 0 || o.cmvelreftype == 1, VRGPU_EINVAL, "Kinetic_reference_frame_type must be 0 or 1");
    VR_REQUIRE(o.bgpot == 0 || o.bgpot == 1, VRGPU_EINVAL, "Keep_background_potential must be 0 or 1");
    // Keep_background_potential = 0 in the large-group loop (unbind.cxx:577-606): pair terms are taken off only when
    // nEplus < 2 ln n, and a pass happens only when nEplus >= maxallowedunboundfrac * n (FillUnboundArrays), so with
    // maxallowedunboundfrac * n >= 2 ln n for every n >= ompunbindnum (the defaults: 25 >= 13.8) that branch never runs;
    // the other branch re-evaluates Potential() over the still complete member list, i.e. the values it already has.
    if (o.bgpot == 0) {
        const double a = o.maxallowedunboundfrac, Tl = (double)o.large_threshold; // a n - 2 ln n grows with n from n = 2 / a on
        VR_REQUIRE(a > 0 && Tl >= 2.0 / a && a * Tl >= 2.0 * std::log(Tl), VRGPU_EUNSUPPORTED,
                   "Keep_background_potential=0 needs maxallowedunboundfrac * ompunbindnum >= 2 ln(ompunbindnum) (pair-term branch of the large-group loop)");
    }
    VR_REQUIRE(o.large_threshold >= 1 && o.large_threshold <= UB_CAP, VRGPU_EUNSUPPORTED, "large_threshold must be in [1, 2048]");
    VR_REQUIRE(o.unbindtype == 0 || o.unbindtype == 1, VRGPU_EINVAL, "unbindtype must be 0 or 1");
    VR_REQUIRE(P.has_vel, VRGPU_EINVAL, "unbinding needs velocities");
    const int T = 256;
    GroupSet gs;
    { KScope k(ctx, "unbind.regroup"); build_groupset(ctx, pfof_dev, P.n, ngroups, gs); }
    if (gs.ntot == 0) return;
    { KScope k(ctx, "unbind.gather"); gather_groupset(ctx, P, gs, true, true); }
    const uint32_t ntot = (uint32_t)gs.ntot, G = (uint32_t)ngroups;
    DBuf<double> phi(ctx, ntot);
    compute_potentials(ctx, gs, o.G, o.eps, o.theta, o.bucket, o.pp_threshold, phi.p);

    StageTimer t(ctx);
    const size_t G1 = (size_t)G + 1;
    DBuf<double> E(ctx, ntot), cmvel(ctx, 3 * G1), gmass(ctx, G1);
    DBuf<uint32_t> ord(ctx, ntot), ning(ctx, G1), nunb(ctx, G1), nefrac(ctx, G1), nrem(ctx, G1);
    DBuf<unsigned char> alive(ctx, ntot), marg(ctx, ntot);
    DBuf<unsigned long long> maxE(ctx, G1);
    DBuf<int> status(ctx, G1);
    DBuf<unsigned int> chg(ctx, 1);
    VR_CUDA(cudaMemsetAsync(nunb.p, 0, sizeof(uint32_t) * G1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(nefrac.p, 0, sizeof(uint32_t) * G1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(nrem.p, 0, sizeof(uint32_t) * G1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(ning.p, 0, sizeof(uint32_t) * G1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(maxE.p, 0, sizeof(unsigned long long) * G1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(status.p, 0, sizeof(int) * G1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(chg.p, 0, sizeof(unsigned int), ctx->stream));
    UbArrays A{};
    A.vx = gs.vx.p; A.vy = gs.vy.p; A.vz = gs.vz.p; A.m = gs.m.p; A.phi = phi.p;
    A.gid = gs.gid.p; A.off = gs.off.p;
    A.E = E.p; A.ord = ord.p; A.alive = alive.p; A.marg = marg.p;
    A.cmvel = cmvel.p; A.gmass = gmass.p; A.ning = ning.p; A.nunb = nunb.p; A.nefrac = nefrac.p; A.maxE = maxE.p; A.nrem = nrem.p;
    A.status = status.p; A.changed = chg.p;
    UbParams Pm{};
    Pm.Eratio = o.Eratio; Pm.minEfrac = o.minEfrac; Pm.maxunbindfrac = o.maxunbindfrac;
    Pm.maxunboundfrac = o.maxunboundfracforiterativeunbind; Pm.maxallowed = o.maxallowedunboundfrac;
    Pm.marginal_tol = o.marginal_tol; Pm.unbindtype = o.unbindtype; Pm.min_size = o.min_size; Pm.large_threshold = o.large_threshold;
    Pm.fixed_frame = o.fracpotref != 1.0 ? 1 : 0;
    const unsigned warp_grid = std::max(1u, std::min<unsigned>(grid_for((size_t)G * 32, T), (unsigned)ctx->sm_count * 16u));

    { KScope k(ctx, "unbind.cm_frame"); VR_LAUNCH(ctx, k_init_slots, grid_for(ntot, T), T, 0, A, ntot);
    if (!Pm.fixed_frame) {
        VR_LAUNCH(ctx, k_group_cm_big, G, 1024, 0, A);
        VR_LAUNCH(ctx, k_group_cm, warp_grid, T, 0, A, G);
    } else {
        // radius order inside every group: stable sort by radius, then by group id
        DBuf<double> centre(ctx, 3 * G1);
        DBuf<uint64_t> keys(ctx, ntot), keys_tmp(ctx, ntot);
        DBuf<uint32_t> vals(ctx, ntot), vals_tmp(ctx, ntot);
        VR_LAUNCH(ctx, k_frame_centre, G, FR_T, 0, gs.x.p, gs.y.p, gs.z.p, gs.m.p, phi.p, gs.off.p, o.cmvelreftype == 1 ? 1 : 0, centre.p, gmass.p);
        VR_LAUNCH(ctx, k_frame_radius, grid_for(ntot, T), T, 0, gs.x.p, gs.y.p, gs.z.p, gs.gid.p, centre.p, ntot, keys.p, vals.p);
        radix_sort_pairs(ctx, keys.p, vals.p, keys_tmp.p, vals_tmp.p, ntot, 64);
        VR_LAUNCH(ctx, k_frame_gidkeys, grid_for(ntot, T), T, 0, gs.gid.p, vals.p, ntot, keys.p);
        int gbits = 1;
        while ((1ull << gbits) <= (unsigned long long)G) gbits++;
        radix_sort_pairs(ctx, keys.p, vals.p, keys_tmp.p, vals_tmp.p, ntot, gbits);
        VR_LAUNCH(ctx, k_frame_cmvel, G, FR_T, 0, A, vals.p, o.fracpotref, (long long)o.npotref);
    } }

    // ---- phase A ----
    {
        KScope ka(ctx, "unbind.large_groups");
        DBuf<uint32_t> lflag(ctx, ntot), lexcl(ctx, ntot), ltotal(ctx, 1);
        VR_LAUNCH(ctx, k_large_energy, grid_for(ntot, T), T, 0, A, Pm, ntot, lflag.p);
        exclusive_scan_u32(ctx, lflag.p, lexcl.p, ntot, ltotal.p);
        uint32_t nl = 0;
        VR_CUDA(cudaMemcpyAsync(&nl, ltotal.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        if (nl > 0) {
            DBuf<uint64_t> keys(ctx, nl), keys_tmp(ctx, nl);
            DBuf<uint32_t> vals(ctx, nl), vals_tmp(ctx, nl), lstart(ctx, G1 + 1);
            VR_LAUNCH(ctx, k_large_keys, grid_for(ntot, T), T, 0, A, lflag.p, lexcl.p, ntot, keys.p, vals.p);
            radix_sort_pairs(ctx, keys.p, vals.p, keys_tmp.p, vals_tmp.p, nl, 64);
            VR_LAUNCH(ctx, k_large_gidkeys, grid_for(nl, T), T, 0, A, vals.p, nl, keys.p);
            int gbits = 1;
            while ((1ull << gbits) <= (unsigned long long)G) gbits++;
            radix_sort_pairs(ctx, keys.p, vals.p, keys_tmp.p, vals_tmp.p, nl, gbits);
            VR_LAUNCH(ctx, k_large_decide, grid_for(G, T), T, 0, A, Pm, G);
            // lstart[g] = position of group g's first member in the compacted gid-sorted list = lexcl[off[g]]
            // (slots of one group are contiguous and a large group is compacted whole)
            VR_LAUNCH(ctx, k_large_lstart, grid_for(G1, T), T, 0, gs.off.p, lexcl.p, G, ntot, lstart.p);
            VR_LAUNCH(ctx, k_large_apply, grid_for(nl, T), T, 0, A, vals.p, lstart.p, nl);
            VR_LAUNCH(ctx, k_large_finish, warp_grid, T, 0, A, Pm, G);
        }
    }
    // ---- phase B ----
    {
        KScope k(ctx, "unbind.small_groups");
        UbBg Bg{gs.x.p, gs.y.p, gs.z.p, phi.p, o.G, o.eps * o.eps};
        if (o.bgpot == 0) VR_LAUNCH(ctx, k_unbind_small<true>, G, UB_T, 0, A, Pm, G, Bg);
        else VR_LAUNCH(ctx, k_unbind_small<false>, G, UB_T, 0, A, Pm, G, Bg);
    }

    // ---- results ----
    DBuf<uint32_t> newsize(ctx, G1), iota(ctx, G1), gid_of(ctx, G1);
    DBuf<unsigned long long> minpid(ctx, G1), nmarg(ctx, 1);
    VR_CUDA(cudaMemsetAsync(newsize.p, 0, sizeof(uint32_t) * G1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(minpid.p, 0xff, sizeof(unsigned long long) * G1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(nmarg.p, 0, sizeof(unsigned long long), ctx->stream));
    VR_LAUNCH(ctx, k_new_group_stats, grid_for(ntot, T), T, 0, A, gs.idx.p, P.has_pid ? P.pid.p : (const long long *)nullptr, ntot, newsize.p,
              minpid.p, nmarg.p);
    unsigned int hchg = 0;
    unsigned long long hmarg = 0;
    VR_CUDA(cudaMemcpyAsync(&hchg, chg.p, sizeof(hchg), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(&hmarg, nmarg.p, sizeof(hmarg), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    n_marginal = (long long)hmarg;
    changed = hchg ? 1 : 0;
    if (changed) {
        VR_LAUNCH(ctx, k_iota_u32, grid_for(G1, T), T, 0, iota.p, G + 1);
        uint32_t ng = order_groups(ctx, iota.p, newsize.p, minpid.p, G + 1, 1u, gid_of.p);
        VR_LAUNCH(ctx, k_write_unbound, grid_for(ntot, T), T, 0, A, gs.idx.p, gid_of.p, ntot, pfof_dev);
        ngroups = ng;
    }
    ctx->stats.unbind_ms += t.stop();
}

// groups handed over in CSR form by the host (vrgpu_potential, vrgpu_tree_order) -> group-major layout
static bool groupset_from_csr(vrgpu_ctx *ctx, const DeviceParticles &P, long long ngroups, const long long *offsets_host,
                              const long long *index_host, GroupSet &gs)
{
    const long long ntot = offsets_host[ngroups];
    if (ntot == 0) return false;
    VR_REQUIRE(ntot < 0xfffffff0LL && ngroups < 0x7fffffffLL, VRGPU_EINVAL, "CSR too large");
    VR_REQUIRE(offsets_host[0] == 0, VRGPU_EINVAL, "offsets[0] must be 0");
    std::vector<uint32_t> off((size_t)ngroups + 2), idx((size_t)ntot);
    off[0] = 0;
    for (long long g = 0; g <= ngroups; g++) {
        VR_REQUIRE(offsets_host[g] >= 0 && (g == 0 || offsets_host[g] >= offsets_host[g - 1]), VRGPU_EINVAL, "offsets must be non-decreasing");
        off[(size_t)g + 1] = (uint32_t)offsets_host[g];
    }
    for (long long k = 0; k < ntot; k++) {
        VR_REQUIRE(index_host[k] >= 0 && (size_t)index_host[k] < P.n, VRGPU_EINVAL, "index out of range");
        idx[(size_t)k] = (uint32_t)index_host[k];
    }
    gs.ntot = (size_t)ntot;
    gs.ngroups = ngroups;
    gs.off.alloc(ctx, off.size());
    gs.idx.alloc(ctx, idx.size());
    gs.gid.alloc(ctx, idx.size());
    VR_CUDA(cudaMemcpyAsync(gs.off.p, off.data(), off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(gs.idx.p, idx.data(), idx.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    VR_LAUNCH(ctx, k_csr_gid, std::max(1u, std::min<unsigned>(grid_for((size_t)ngroups * 32, 256), (unsigned)ctx->sm_count * 16u)), 256, 0,
              gs.off.p, (uint32_t)ngroups, gs.gid.p);
    return true;
}

// vrgpu_potential
void potential_csr_host(vrgpu_ctx *ctx, const DeviceParticles &P, long long ngroups, const long long *offsets_host,
                        const long long *index_host, double G, double eps, double theta, int bucket, long long pp_threshold,
                        double *phi_out_host)
{
    GroupSet gs;
    if (!groupset_from_csr(ctx, P, ngroups, offsets_host, index_host, gs)) return;
    const long long ntot = (long long)gs.ntot;
    gather_groupset(ctx, P, gs, false, true);
    DBuf<double> phi(ctx, (size_t)ntot);
    compute_potentials(ctx, gs, G, eps, theta, bucket, pp_threshold, phi.p);
    VR_CUDA(cudaMemcpyAsync(phi_out_host, phi.p, (size_t)ntot * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->stats.d2h_bytes += ntot * 8;
}

// vrgpu_tree_order: the arrangement BuildNodes leaves (every group a tree of its own)
void tree_order_csr_host(vrgpu_ctx *ctx, const DeviceParticles &P, long long ngroups, const long long *offsets_host,
                         const long long *index_host, int bucket, long long *perm_out_host)
{
    GroupSet gs;
    if (!groupset_from_csr(ctx, P, ngroups, offsets_host, index_host, gs)) return;
    gather_groupset(ctx, P, gs, false, false);
    DBuf<long long> out(ctx, gs.ntot);
    tree_order_device(ctx, gs, bucket, out.p);
    VR_CUDA(cudaMemcpyAsync(perm_out_host, out.p, gs.ntot * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->stats.d2h_bytes += (long long)gs.ntot * 8;
}

} // namespace vrgpu
