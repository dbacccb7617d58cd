// velociraptor-stf_b200/csrc/sort.cu -- device primitives: stable LSD radix sort of (key64,val32) pairs and
// exclusive prefix sums.  Hand-written for sm_100a; HBM-bound integer work (no tensor cores).
//
// Radix pass = 3 launches over tiles of RS_TILE elements:
//   rs_hist     per-tile digit histogram (warp-aggregated shared atomics)  -> table[digit][tile]
//   rs_scan     one CTA per digit: exclusive scan of its row, row total    -> table (in place), totals[digit]
//   rs_scatter  re-read tile, stable in-tile ranks via __match_any_sync, order the tile by digit in shared memory,
//               write each digit's run to its final position with coalesced stores
// Algorithmic traffic per pass: 8 B (hist) + 12 B read + 12 B write per element.
#include "common.cuh"

namespace vrgpu {

constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 8;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS; // 4096
constexpr int RS_BITS = 8;
constexpr int RS_BINS = 1 << RS_BITS;
constexpr int RS_WARPS = RS_THREADS / 32;

__global__ void __launch_bounds__(RS_THREADS) rs_hist(const uint64_t *__restrict__ keys, size_t n, int shift,
                                                      uint32_t *__restrict__ table, uint32_t ntiles)
{
    __shared__ uint32_t h[RS_BINS];
    h[threadIdx.x] = 0;
    __syncthreads();
    const size_t base = (size_t)blockIdx.x * RS_TILE;
    const unsigned lane = threadIdx.x & 31;
#pragma unroll 4
    for (int r = 0; r < RS_ITEMS; r++) {
        size_t i = base + (size_t)r * RS_THREADS + threadIdx.x;
        unsigned digit = RS_BINS; // sentinel for out-of-range lanes
        if (i < n) digit = (unsigned)((keys[i] >> shift) & (RS_BINS - 1));
        unsigned peers = __match_any_sync(0xffffffffu, digit);
        if (digit < RS_BINS && lane == (unsigned)(__ffs(peers) - 1)) atomicAdd(&h[digit], __popc(peers));
    }
    __syncthreads();
    table[(size_t)threadIdx.x * ntiles + blockIdx.x] = h[threadIdx.x];
}

// one CTA per digit: exclusive scan of table[digit][0..ntiles) in place, totals[digit] = row sum
__global__ void __launch_bounds__(256) rs_scan(uint32_t *__restrict__ table, uint32_t ntiles, uint32_t *__restrict__ totals)
{
    __shared__ uint32_t wsum[8];
    __shared__ uint32_t carry_s;
    uint32_t *row = table + (size_t)blockIdx.x * ntiles;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < ntiles; base += 256) {
        uint32_t i = base + threadIdx.x;
        uint32_t v = i < ntiles ? row[i] : 0;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= (unsigned)o) x += y;
        }
        if (lane == 31) wsum[warp] = x;
        __syncthreads();
        uint32_t woff = 0;
        for (unsigned w = 0; w < warp; w++) woff += wsum[w];
        uint32_t carry = carry_s;
        if (i < ntiles) row[i] = carry + woff + x - v;
        __syncthreads();
        if (threadIdx.x == 255) carry_s = carry + woff + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry_s;
}

// The tile is first ordered by digit in shared memory (stable: warp-major, then load order), then written out with
// consecutive threads on consecutive addresses: a digit's run of a tile (16 elements on average) leaves as whole sectors
// instead of one 8-byte + one 4-byte store per element to 256 different places.
constexpr size_t RS_SMEM = (size_t)RS_TILE * 12 + (size_t)RS_WARPS * RS_BINS * 4 + 2 * RS_BINS * 4 + 64;

__global__ void __launch_bounds__(RS_THREADS) rs_scatter(const uint64_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                                                         uint64_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, size_t n,
                                                         int shift, const uint32_t *__restrict__ table, uint32_t ntiles,
                                                         const uint32_t *__restrict__ totals)
{
    extern __shared__ __align__(16) unsigned char rs_smem[];
    uint64_t *skey = reinterpret_cast<uint64_t *>(rs_smem);
    uint32_t *sval = reinterpret_cast<uint32_t *>(skey + RS_TILE);
    uint32_t(*whist)[RS_BINS] = reinterpret_cast<uint32_t(*)[RS_BINS]>(sval + RS_TILE);
    uint32_t *dbase = reinterpret_cast<uint32_t *>(whist + RS_WARPS); // global start of digit d
    uint32_t *gadj = dbase + RS_BINS;                                  // global position of tile-local index 0 of digit d
    uint32_t *wsum = gadj + RS_BINS;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int w = 0; w < RS_WARPS; w++) whist[w][threadIdx.x] = 0;
    // exclusive scan of the 256 digit totals -> dbase
    {
        uint32_t v = totals[threadIdx.x], x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= (unsigned)o) x += y;
        }
        if (lane == 31) wsum[warp] = x;
        __syncthreads();
        uint32_t woff = 0;
        for (unsigned w = 0; w < warp; w++) woff += wsum[w];
        dbase[threadIdx.x] = woff + x - v;
    }
    __syncthreads();

    uint64_t k[RS_ITEMS];
    uint32_t v[RS_ITEMS];
    uint16_t rank[RS_ITEMS];
    const size_t tbase = (size_t)blockIdx.x * RS_TILE;
    const size_t wbase = tbase + (size_t)warp * (RS_ITEMS * 32);
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++) {
        size_t i = wbase + (size_t)r * 32 + lane;
        unsigned digit = RS_BINS;
        if (i < n) {
            k[r] = keys_in[i];
            v[r] = vals_in[i];
            digit = (unsigned)((k[r] >> shift) & (RS_BINS - 1));
        }
        unsigned peers = __match_any_sync(0xffffffffu, digit);
        uint32_t before = 0;
        if (digit < RS_BINS) before = whist[warp][digit];
        __syncwarp();
        if (digit < RS_BINS && lane == (unsigned)(__ffs(peers) - 1)) whist[warp][digit] = before + __popc(peers);
        __syncwarp();
        rank[r] = (uint16_t)(before + __popc(peers & lt));
    }
    __syncthreads();
    {
        // thread d owns digit d: tile-local start of the digit (exclusive scan over digits), per-warp offsets inside it
        const unsigned d = threadIdx.x;
        uint32_t tot = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) tot += whist[w][d];
        uint32_t x = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= (unsigned)o) x += y;
        }
        __syncthreads(); // wsum is reused
        if (lane == 31) wsum[warp] = x;
        __syncthreads();
        uint32_t woff = 0;
        for (unsigned w = 0; w < warp; w++) woff += wsum[w];
        const uint32_t tstart = woff + x - tot;
        uint32_t running = tstart;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            uint32_t t = whist[w][d];
            whist[w][d] = running;
            running += t;
        }
        gadj[d] = dbase[d] + table[(size_t)d * ntiles + blockIdx.x] - tstart;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; r++) {
        size_t i = wbase + (size_t)r * 32 + lane;
        if (i < n) {
            unsigned digit = (unsigned)((k[r] >> shift) & (RS_BINS - 1));
            uint32_t lp = whist[warp][digit] + rank[r];
            skey[lp] = k[r];
            sval[lp] = v[r];
        }
    }
    __syncthreads();
    const uint32_t cnt = (uint32_t)min((size_t)RS_TILE, n - tbase);
    for (uint32_t idx = threadIdx.x; idx < cnt; idx += RS_THREADS) {
        const uint64_t key = skey[idx];
        const uint32_t pos = gadj[(unsigned)((key >> shift) & (RS_BINS - 1))] + idx;
        keys_out[pos] = key;
        vals_out[pos] = sval[idx];
    }
}

void radix_sort_pairs(vrgpu_ctx *ctx, uint64_t *keys, uint32_t *vals, uint64_t *keys_tmp, uint32_t *vals_tmp, size_t n,
                      int nbits)
{
    if (n == 0 || nbits <= 0) return;
    VR_REQUIRE(n < (size_t)0xffffffffu, VRGPU_EINVAL, "radix_sort_pairs: n must fit in 32 bits");
    const uint32_t ntiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
    DBuf<uint32_t> table(ctx, (size_t)RS_BINS * ntiles);
    DBuf<uint32_t> totals(ctx, RS_BINS);
    uint64_t *kin = keys, *kout = keys_tmp;
    uint32_t *vin = vals, *vout = vals_tmp;
    int npass = (nbits + RS_BITS - 1) / RS_BITS;
    VR_CUDA(cudaFuncSetAttribute(rs_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RS_SMEM));
    for (int p = 0; p < npass; p++) {
        int shift = p * RS_BITS;
        VR_LAUNCH(ctx, rs_hist, ntiles, RS_THREADS, 0, kin, n, shift, table.p, ntiles);
        VR_LAUNCH(ctx, rs_scan, RS_BINS, 256, 0, table.p, ntiles, totals.p);
        VR_LAUNCH(ctx, rs_scatter, ntiles, RS_THREADS, RS_SMEM, kin, vin, kout, vout, n, shift, table.p, ntiles, totals.p);
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    if (kin != keys) {
        VR_CUDA(cudaMemcpyAsync(keys, kin, n * sizeof(uint64_t), cudaMemcpyDeviceToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(vals, vin, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    }
}

// ------------------------------------------------------------------------------------------------
// exclusive scan (reduce / recurse / apply), tiles of 1024 threads x 4 items
// ------------------------------------------------------------------------------------------------
constexpr int SC_THREADS = 1024;
constexpr int SC_ITEMS = 4;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;

template <typename T> __device__ __forceinline__ T block_excl_scan(T v, T *total)
{
    __shared__ T wsum[32];
    __shared__ T tot;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    if (warp == 0) {
        T w = lane < (blockDim.x >> 5) ? wsum[lane] : T(0);
        T s = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            T y = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= (unsigned)o) s += y;
        }
        wsum[lane] = s - w;
        if (lane == 31) tot = s;
    }
    __syncthreads();
    T r = wsum[warp] + x - v;
    if (total) *total = tot;
    __syncthreads();
    return r;
}

template <typename T> __global__ void __launch_bounds__(SC_THREADS) sc_reduce(const T *__restrict__ in, size_t n, T *__restrict__ bsum)
{
    size_t base = (size_t)blockIdx.x * SC_TILE + (size_t)threadIdx.x * SC_ITEMS;
    T s = 0;
#pragma unroll
    for (int k = 0; k < SC_ITEMS; k++) if (base + k < n) s += in[base + k];
    T tot;
    block_excl_scan<T>(s, &tot);
    if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}

template <typename T>
__global__ void __launch_bounds__(SC_THREADS) sc_apply(const T *__restrict__ in, T *__restrict__ out, size_t n, const T *__restrict__ boff,
                                                       T *__restrict__ total_dev)
{
    size_t base = (size_t)blockIdx.x * SC_TILE + (size_t)threadIdx.x * SC_ITEMS;
    T v[SC_ITEMS];
    T s = 0;
#pragma unroll
    for (int k = 0; k < SC_ITEMS; k++) { v[k] = base + k < n ? in[base + k] : T(0); s += v[k]; }
    T tot;
    T off = block_excl_scan<T>(s, &tot) + (boff ? boff[blockIdx.x] : T(0));
#pragma unroll
    for (int k = 0; k < SC_ITEMS; k++) {
        if (base + k < n) out[base + k] = off;
        off += v[k];
    }
    if (total_dev && blockIdx.x == gridDim.x - 1 && threadIdx.x == SC_THREADS - 1) *total_dev = off;
}

template <typename T> static void exclusive_scan_t(vrgpu_ctx *ctx, const T *in, T *out, size_t n, T *total_dev)
{
    if (n == 0) {
        if (total_dev) VR_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(T), ctx->stream));
        return;
    }
    size_t nb = (n + SC_TILE - 1) / SC_TILE;
    if (nb == 1) {
        VR_LAUNCH(ctx, sc_apply<T>, 1, SC_THREADS, 0, in, out, n, (const T *)nullptr, total_dev);
        return;
    }
    DBuf<T> bsum(ctx, nb);
    VR_LAUNCH(ctx, sc_reduce<T>, (unsigned)nb, SC_THREADS, 0, in, n, bsum.p);
    exclusive_scan_t<T>(ctx, bsum.p, bsum.p, nb, nullptr);
    VR_LAUNCH(ctx, sc_apply<T>, (unsigned)nb, SC_THREADS, 0, in, out, n, (const T *)bsum.p, total_dev);
}

void exclusive_scan_u32(vrgpu_ctx *ctx, const uint32_t *in, uint32_t *out, size_t n, uint32_t *total_dev)
{
    exclusive_scan_t<uint32_t>(ctx, in, out, n, total_dev);
}
void exclusive_scan_u64(vrgpu_ctx *ctx, const unsigned long long *in, unsigned long long *out, size_t n,
                        unsigned long long *total_dev)
{
    exclusive_scan_t<unsigned long long>(ctx, in, out, n, total_dev);
}

} // namespace vrgpu
