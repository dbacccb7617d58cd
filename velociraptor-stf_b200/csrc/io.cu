// velociraptor-stf_b200/csrc/io.cu -- host <-> device marshalling of the caller's strided particle view.
//
// The reference holds particles as a 96-byte AoS (NBody::Particle, NBodylib/src/NBody/Particle.h:75-109); the
// device wants packed columns.  Chunks of the raw records are copied with cudaMemcpyAsync (copy stream, double
// buffered) into device staging buffers and unpacked there by k_unpack on the context stream.
#include "fof.cuh"

#include <algorithm>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>
#include <string>

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

namespace vrgpu {

struct UnpackArgs {
    const unsigned char *base; // device-visible pointer to record 0 of this chunk
    size_t stride;
    size_t off_pos, off_vel, off_mass, off_pid;
    size_t n;      // records in this chunk
    size_t first;  // global index of record 0
    double *pos, *vel, *mass;
    long long *pid;
};

__global__ void k_unpack(UnpackArgs a)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const unsigned char *rec = a.base + i * a.stride;
    const size_t o = a.first + i;
    if (a.pos) {
        const double *p = reinterpret_cast<const double *>(rec + a.off_pos);
        a.pos[3 * o] = p[0]; a.pos[3 * o + 1] = p[1]; a.pos[3 * o + 2] = p[2];
    }
    if (a.vel) {
        const double *p = reinterpret_cast<const double *>(rec + a.off_vel);
        a.vel[3 * o] = p[0]; a.vel[3 * o + 1] = p[1]; a.vel[3 * o + 2] = p[2];
    }
    if (a.mass) a.mass[o] = *reinterpret_cast<const double *>(rec + a.off_mass);
    if (a.pid) a.pid[o] = *reinterpret_cast<const long long *>(rec + a.off_pid);
}

namespace {
// events borrowed from the context's pool; on the way out (also by exception) the copy stream is drained first, so that no
// staging buffer goes back to the pool -- and no event is reused -- while a copy may still be in flight
struct UploadEvents {
    vrgpu_ctx *ctx;
    std::vector<cudaEvent_t> ev;
    explicit UploadEvents(vrgpu_ctx *c, int n) : ctx(c) { for (int k = 0; k < n; k++) ev.push_back(KScope::get(c)); }
    ~UploadEvents()
    {
        if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
        for (cudaEvent_t e : ev) ctx->evpool.push_back(e);
    }
};

// is the caller's memory page-locked (cudaHostAlloc / cudaHostRegister)?  Pageable memory is what the reference hands over
// (vector<Particle>, new Int_t[n]).
bool host_memory_is_pinned(const void *p)
{
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

// copy the byte window [wlo, wlo + width) of records [0, cnt) into a packed buffer (host threads share the records)
void pack_window(unsigned char *dst, const unsigned char *src, size_t cnt, size_t stride, size_t wlo, size_t width, int nthreads)
{
    auto work = [=](size_t r0, size_t r1) {
        if (width == stride) { std::memcpy(dst + r0 * width, src + r0 * stride, (r1 - r0) * width); return; }
        const size_t nw = width / 8;
#if defined(__SSE2__)
        // 64-byte windows into the 64-byte aligned staging buffer (NBody::Particle: mass, pos, vel, pid): streaming stores, so that
        // the staging lines are not read before they are overwritten -- the packing is bound by the host's memory bandwidth
        if (width == 64 && (reinterpret_cast<uintptr_t>(dst) & 63u) == 0) {
            for (size_t i = r0; i < r1; i++) {
                const __m128i *s16 = reinterpret_cast<const __m128i *>(src + i * stride + wlo);
                __m128i *d16 = reinterpret_cast<__m128i *>(dst + i * 64);
                const __m128i a = _mm_loadu_si128(s16), b = _mm_loadu_si128(s16 + 1), c = _mm_loadu_si128(s16 + 2), d = _mm_loadu_si128(s16 + 3);
                _mm_stream_si128(d16, a); _mm_stream_si128(d16 + 1, b); _mm_stream_si128(d16 + 2, c); _mm_stream_si128(d16 + 3, d);
            }
            _mm_sfence();
            return;
        }
#endif
        for (size_t i = r0; i < r1; i++) {
            const uint64_t *s8 = reinterpret_cast<const uint64_t *>(src + i * stride + wlo);
            uint64_t *d8 = reinterpret_cast<uint64_t *>(dst + i * width);
            for (size_t k = 0; k < nw; k++) d8[k] = s8[k];
        }
    };
    if (nthreads <= 1 || cnt < 65536) { work(0, cnt); return; }
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; t++) th.emplace_back(work, cnt * t / nthreads, cnt * (t + 1) / nthreads);
    work(0, cnt / nthreads);
    for (auto &x : th) x.join();
}
} // namespace

void upload_view(vrgpu_ctx *ctx, const vrgpu_view &v, bool need_vel, bool need_mass, bool need_pid, DeviceParticles &P)
{
    VR_REQUIRE(v.n >= 0, VRGPU_EINVAL, "negative particle count");
    const size_t n = (size_t)v.n;
    P.n = n;
    P.has_vel = need_vel && v.off_vel != VRGPU_NOFIELD;
    P.has_mass = need_mass && v.off_mass != VRGPU_NOFIELD;
    P.has_pid = need_pid && v.off_pid != VRGPU_NOFIELD;
    VR_REQUIRE(!need_vel || P.has_vel, VRGPU_EINVAL, "this call needs velocities (view.off_vel)");
    if (n == 0) return;
    VR_REQUIRE(v.base != nullptr, VRGPU_EINVAL, "view.base is NULL");
    VR_REQUIRE(v.stride % 8 == 0 && v.stride >= 24, VRGPU_EINVAL, "view.stride must be a multiple of 8 and >= 24");
    VR_REQUIRE(((uintptr_t)v.base % 8) == 0, VRGPU_EINVAL, "view.base must be 8-byte aligned");
    // every field used lies inside the record and is 8-byte aligned; [wlo, whi) = the bytes of a record that are read
    size_t wlo = v.off_pos, whi = v.off_pos + 24;
    auto field = [&](bool used, size_t off, size_t size, const char *what) {
        if (!used) return;
        VR_REQUIRE(off % 8 == 0 && off + size <= v.stride, VRGPU_EINVAL, std::string("view field outside the record or misaligned: ") + what);
        wlo = std::min(wlo, off);
        whi = std::max(whi, off + size);
    };
    field(true, v.off_pos, 24, "off_pos");
    field(P.has_vel, v.off_vel, 24, "off_vel");
    field(P.has_mass, v.off_mass, 8, "off_mass");
    field(P.has_pid, v.off_pid, 8, "off_pid");
    P.pos.alloc(ctx, 3 * n);
    if (P.has_vel) P.vel.alloc(ctx, 3 * n);
    if (P.has_mass) P.mass.alloc(ctx, n);
    if (P.has_pid) P.pid.alloc(ctx, n);
    const int T = 256;
    UnpackArgs a{};
    a.pos = P.pos.p;
    a.vel = P.has_vel ? P.vel.p : nullptr;
    a.mass = P.has_mass ? P.mass.p : nullptr;
    a.pid = P.has_pid ? P.pid.p : nullptr;
    if (!ctx->copy_stream) VR_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    KScope kup(ctx, "io.upload");
    // Chunks of records travel on a second stream into two device staging buffers and are unpacked on the context stream while
    // the next chunk is in flight.
    //   page-locked caller memory: whole records, straight from the caller (54 GB/s measured; a strided 2-D DMA of the used
    //     bytes only: 2.5 GB/s; reading the records zero-copy from the kernel: 21 GB/s -- profiles/r01a);
    //   pageable caller memory (what the reference hands over): host threads pack the USED bytes of every record -- the
    //     first 64 of NBody::Particle's 96 -- into page-locked staging while the previous chunk's DMA runs.  Letting the
    //     driver bounce pageable memory itself measured 11 GB/s (profiles/r02a_probe.md).
    const bool pinned = host_memory_is_pinned(v.base);
    const size_t width = pinned ? v.stride : whi - wlo;
    const size_t base_off = pinned ? 0 : wlo;
    a.stride = width;
    a.off_pos = v.off_pos - base_off;
    a.off_vel = P.has_vel ? v.off_vel - base_off : 0;
    a.off_mass = P.has_mass ? v.off_mass - base_off : 0;
    a.off_pid = P.has_pid ? v.off_pid - base_off : 0;
    const size_t chunk_bytes = pinned ? (size_t)64 << 20 : vrgpu_ctx::UP_HOST_BYTES;
    size_t per = chunk_bytes / width;
    if (per < 1) per = 1;
    if (per > n) per = n;
    constexpr int NH = 3;
    if (!pinned && !ctx->up_host[0])
        for (int k = 0; k < NH; k++) VR_CUDA(cudaHostAlloc(&ctx->up_host[k], vrgpu_ctx::UP_HOST_BYTES, cudaHostAllocDefault));
    const int nthreads = (int)std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    DBuf<unsigned char> stage[2];
    for (int b = 0; b < 2; b++) stage[b].alloc(ctx, per * width);
    UploadEvents E(ctx, 4 + NH); // ready[2], freed[2], hdone[NH]
    cudaEvent_t *ready = E.ev.data(), *freed = ready + 2, *hdone = ready + 4;
    // the staging buffers may still be in use by earlier work of the context stream
    VR_CUDA(cudaEventRecord(freed[0], ctx->stream));
    VR_CUDA(cudaEventRecord(freed[1], ctx->stream));
    int b = 0;
    size_t c = 0;
    for (size_t first = 0; first < n; first += per, b ^= 1, c++) {
        const size_t cnt = n - first < per ? n - first : per;
        const unsigned char *src = static_cast<const unsigned char *>(v.base) + first * v.stride;
        if (!pinned) {
            const int hb = (int)(c % NH);
            if (c >= (size_t)NH) VR_CUDA(cudaEventSynchronize(hdone[hb])); // the DMA that read this buffer last has finished
            pack_window(static_cast<unsigned char *>(ctx->up_host[hb]), src, cnt, v.stride, wlo, width, nthreads);
            src = static_cast<const unsigned char *>(ctx->up_host[hb]);
        }
        VR_CUDA(cudaStreamWaitEvent(ctx->copy_stream, freed[b], 0));
        VR_CUDA(cudaMemcpyAsync(stage[b].p, src, cnt * width, cudaMemcpyHostToDevice, ctx->copy_stream));
        if (!pinned) VR_CUDA(cudaEventRecord(hdone[c % NH], ctx->copy_stream));
        VR_CUDA(cudaEventRecord(ready[b], ctx->copy_stream));
        VR_CUDA(cudaStreamWaitEvent(ctx->stream, ready[b], 0));
        a.base = stage[b].p;
        a.n = cnt;
        a.first = first;
        VR_LAUNCH(ctx, k_unpack, grid_for(cnt, T), T, 0, a);
        VR_CUDA(cudaEventRecord(freed[b], ctx->stream));
        ctx->stats.h2d_bytes += (long long)(cnt * width);
    }
}

// Results go back in 32 MB chunks through two page-locked staging buffers: the DMA of chunk k+1 runs while host threads
// consume chunk k (copy / scatter it into the caller's pageable, usually never touched memory -- `new Int_t[n]`,
// vector<Particle>).  A plain cudaMemcpy into such memory measured 4.9 GB/s (profiles/r01f).
void staged_download(vrgpu_ctx *ctx, const void *dev, size_t bytes, size_t granule,
                     const std::function<void(const unsigned char *chunk, size_t offset, size_t len, int part, int nparts)> &consume)
{
    if (bytes == 0) return;
    size_t chunk = (size_t)32 << 20;
    chunk -= chunk % granule;
    if (!ctx->stage_host[0]) {
        VR_CUDA(cudaHostAlloc(&ctx->stage_host[0], (size_t)32 << 20, cudaHostAllocDefault));
        VR_CUDA(cudaHostAlloc(&ctx->stage_host[1], (size_t)32 << 20, cudaHostAllocDefault));
    }
    cudaEvent_t done[2] = {KScope::get(ctx), KScope::get(ctx)};
    const unsigned char *src = static_cast<const unsigned char *>(dev);
    const size_t nchunks = (bytes + chunk - 1) / chunk;
    auto issue = [&](size_t c) {
        const size_t o = c * chunk, len = std::min(chunk, bytes - o);
        VR_CUDA(cudaMemcpyAsync(ctx->stage_host[c & 1], src + o, len, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaEventRecord(done[c & 1], ctx->stream));
    };
    issue(0);
    for (size_t c = 0; c < nchunks; c++) {
        VR_CUDA(cudaEventSynchronize(done[c & 1]));
        if (c + 1 < nchunks) issue(c + 1);
        const size_t o = c * chunk, len = std::min(chunk, bytes - o);
        const unsigned char *from = static_cast<const unsigned char *>(ctx->stage_host[c & 1]);
        const int nt = len > ((size_t)1 << 20) ? 8 : 1; // 8 threads measured best (first-touch page faults dominate)
        std::thread th[8];
        for (int t = 1; t < nt; t++) th[t] = std::thread([&, t] { consume(from, o, len, t, nt); });
        consume(from, o, len, 0, nt);
        for (int t = 1; t < nt; t++) th[t].join();
    }
    ctx->evpool.push_back(done[0]);
    ctx->evpool.push_back(done[1]);
    ctx->stats.d2h_bytes += (long long)bytes;
}

void download_i64(vrgpu_ctx *ctx, const long long *dev, long long *host, size_t n)
{
    if (n == 0) return;
    KScope kdn(ctx, "io.download");
    const size_t bytes = n * sizeof(long long);
    cudaPointerAttributes at{};
    const bool locked = cudaPointerGetAttributes(&at, host) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    if (locked || bytes < ((size_t)4 << 20)) {
        VR_CUDA(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->stats.d2h_bytes += (long long)bytes;
        return;
    }
    unsigned char *dst = reinterpret_cast<unsigned char *>(host);
    staged_download(ctx, dev, bytes, 4096, [dst](const unsigned char *chunk, size_t offset, size_t len, int part, int nparts) {
        const size_t per = ((len / nparts) + 4095) & ~(size_t)4095;
        const size_t a0 = std::min(len, (size_t)part * per), a1 = std::min(len, (size_t)(part + 1) * per);
        if (a1 > a0) memcpy(dst + offset + a0, chunk + a0, a1 - a0);
    });
}

} // namespace vrgpu
