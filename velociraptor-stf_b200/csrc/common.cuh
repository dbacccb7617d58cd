// velociraptor-stf_b200/csrc/common.cuh -- shared host/device plumbing for libvrgpu (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/vrgpu.h"

namespace vrgpu {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

#define VR_CUDA(call)                                                                                   \
    do {                                                                                                \
        cudaError_t e__ = (call);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            throw ::vrgpu::Error(VRGPU_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e__) +     \
                                                  " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"); \
    } while (0)

#define VR_REQUIRE(cond, code, msg)                                   \
    do {                                                              \
        if (!(cond)) throw ::vrgpu::Error((code), std::string(msg));  \
    } while (0)

// Caching device allocator: cudaMalloc/cudaFree synchronise the device, so freed blocks are kept and
// re-used by size class for the lifetime of the context.
class Pool {
  public:
    void *alloc(size_t bytes)
    {
        if (bytes == 0) bytes = 1;
        size_t cls = size_class(bytes);
        auto it = free_.lower_bound(cls);
        if (it != free_.end() && it->first <= 2 * cls) {
            void *p = it->second;
            live_[p] = it->first;
            free_.erase(it);
            return p;
        }
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, cls);
        if (e != cudaSuccess) {
            release_cached();
            e = cudaMalloc(&p, cls);
            if (e != cudaSuccess) {
                cudaGetLastError();
                throw Error(VRGPU_ENOMEM, "cudaMalloc of " + std::to_string(cls) + " bytes failed: " + cudaGetErrorString(e));
            }
        }
        live_[p] = cls;
        total_ += cls;
        return p;
    }
    void free(void *p)
    {
        if (!p) return;
        auto it = live_.find(p);
        if (it == live_.end()) return;
        free_.insert({it->second, p});
        live_.erase(it);
    }
    void release_cached()
    {
        for (auto &kv : free_) { cudaFree(kv.second); total_ -= kv.first; }
        free_.clear();
    }
    void release_all()
    {
        release_cached();
        for (auto &kv : live_) cudaFree(kv.first);
        live_.clear();
        total_ = 0;
    }
    size_t total() const { return total_; }

  private:
    static size_t size_class(size_t b)
    {
        const size_t small = 1 << 20;
        if (b <= small) { size_t c = 512; while (c < b) c <<= 1; return c; }
        const size_t g = 2 << 20; // 2 MiB granules
        return (b + g - 1) / g * g;
    }
    std::multimap<size_t, void *> free_;
    std::map<void *, size_t> live_;
    size_t total_ = 0;
};

} // namespace vrgpu

namespace vrgpu {
struct ResidentState; // api.cu: pristine device copy of the caller's particles + last result
struct MultiGpu;      // multigpu.cu: NCCL communicator + slab decomposition state
struct KTimeRec { double ms = 0; long long count = 0; };
struct KPending { const char *name; cudaEvent_t a, b; };
} // namespace vrgpu

struct vrgpu_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr; // host->device staging copies, overlapped with the unpack kernel
    cudaStream_t aux_stream[4] = {nullptr, nullptr, nullptr, nullptr}; // side streams: independent per-group work beside the main stream (fof6d dispersion)
    void *stage_host[2] = {nullptr, nullptr}; // page-locked staging for results that go to pageable caller memory
    static constexpr size_t UP_HOST_BYTES = (size_t)32 << 20;
    void *up_host[3] = {nullptr, nullptr, nullptr}; // page-locked staging for particle records that come from pageable caller memory
    std::string err;
    vrgpu::Pool pool;
    vrgpu_stats stats{};
    // deferred per-kernel timers (CUDA events on `stream`, resolved when the call ends)
    std::vector<vrgpu::KPending> kpending;
    std::vector<cudaEvent_t> evpool;
    std::map<std::string, vrgpu::KTimeRec> ktimes;
    std::string ktimes_text;
    bool ktiming = true;
    // pinned bounce buffers for pageable host memory
    void *pinned = nullptr;
    size_t pinned_bytes = 0;
    vrgpu::ResidentState *resident = nullptr;
    vrgpu::MultiGpu *mg = nullptr;
};

namespace vrgpu {

// RAII device buffer from the context pool
template <typename T> struct DBuf {
    vrgpu_ctx *ctx = nullptr;
    T *p = nullptr;
    size_t n = 0;
    DBuf() = default;
    DBuf(vrgpu_ctx *c, size_t count) : ctx(c), n(count) { p = static_cast<T *>(c->pool.alloc(count * sizeof(T))); }
    DBuf(const DBuf &) = delete;
    DBuf &operator=(const DBuf &) = delete;
    DBuf(DBuf &&o) noexcept : ctx(o.ctx), p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DBuf &operator=(DBuf &&o) noexcept
    {
        if (this != &o) { reset(); ctx = o.ctx; p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
        return *this;
    }
    ~DBuf() { reset(); }
    void reset()
    {
        if (p && ctx) ctx->pool.free(p);
        p = nullptr;
        n = 0;
    }
    void alloc(vrgpu_ctx *c, size_t count)
    {
        reset();
        ctx = c;
        n = count;
        p = static_cast<T *>(c->pool.alloc(count * sizeof(T)));
    }
    operator T *() const { return p; }
    T *get() const { return p; }
    size_t bytes() const { return n * sizeof(T); }
};

inline unsigned grid_for(size_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

#define VR_LAUNCH(ctx, kernel, grid, block, smem, ...)                                   \
    do {                                                                                 \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                 \
        (ctx)->stats.kernel_launches++;                                                  \
        VR_CUDA(cudaGetLastError());                                                     \
    } while (0)

// Deferred timer: brackets a kernel (or a short kernel sequence) with two events on the context stream without
// synchronising; resolve_ktimes() adds the elapsed times to ctx->ktimes[name] once the call has drained.
struct KScope {
    vrgpu_ctx *ctx;
    const char *name;
    cudaEvent_t a = nullptr;
    static cudaEvent_t get(vrgpu_ctx *c)
    {
        if (!c->evpool.empty()) { cudaEvent_t e = c->evpool.back(); c->evpool.pop_back(); return e; }
        cudaEvent_t e = nullptr; // (stays null when the context is in a sticky error state: every user checks)
        if (cudaEventCreate(&e) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        return e;
    }
    KScope(vrgpu_ctx *c, const char *n) : ctx(c), name(n)
    {
        if (!c->ktiming) return;
        a = get(c);
        if (a) cudaEventRecord(a, c->stream);
    }
    ~KScope()
    {
        if (!a) return;
        cudaEvent_t b = get(ctx);
        if (!b) { ctx->evpool.push_back(a); return; }
        cudaEventRecord(b, ctx->stream);
        ctx->kpending.push_back({name, a, b});
    }
};
inline void resolve_ktimes(vrgpu_ctx *ctx)
{
    for (auto &k : ctx->kpending) {
        float ms = 0;
        if (cudaEventSynchronize(k.b) == cudaSuccess && cudaEventElapsedTime(&ms, k.a, k.b) == cudaSuccess) {
            auto &r = ctx->ktimes[k.name];
            r.ms += ms;
            r.count++;
        } else cudaGetLastError();
        ctx->evpool.push_back(k.a);
        ctx->evpool.push_back(k.b);
    }
    ctx->kpending.clear();
}

struct StageTimer {
    vrgpu_ctx *ctx;
    cudaEvent_t a, b;
    StageTimer(vrgpu_ctx *c) : ctx(c)
    {
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        cudaEventRecord(a, c->stream);
    }
    double stop()
    {
        cudaEventRecord(b, ctx->stream);
        cudaEventSynchronize(b);
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        return ms;
    }
    ~StageTimer()
    {
        cudaEventDestroy(a);
        cudaEventDestroy(b);
    }
};

// ---- primitives implemented in sort.cu -------------------------------------------------------------
// Stable LSD radix sort of (key64, val32) pairs on bits [0, nbits).  Result ends in (keys, vals); the
// *_tmp buffers are scratch of the same size.
void radix_sort_pairs(vrgpu_ctx *ctx, uint64_t *keys, uint32_t *vals, uint64_t *keys_tmp, uint32_t *vals_tmp,
                      size_t n, int nbits);
// Exclusive prefix sum of n uint32 (in -> out, may alias); returns nothing, total written to *total_dev (may be null).
void exclusive_scan_u32(vrgpu_ctx *ctx, const uint32_t *in, uint32_t *out, size_t n, uint32_t *total_dev);
// Same for 64-bit counters.
void exclusive_scan_u64(vrgpu_ctx *ctx, const unsigned long long *in, unsigned long long *out, size_t n,
                        unsigned long long *total_dev);

} // namespace vrgpu
