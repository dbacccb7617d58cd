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
    VR_REQUIRE(((uintptr_t)v.base % 8) == 0 && v.off_pos % 8 == 0, VRGPU_EINVAL, "view fields must be 8-byte aligned");
    VR_REQUIRE(v.off_pos + 24 <= v.stride || v.stride == 24, VRGPU_EINVAL, "view.off_pos outside the record");
    P.pos.alloc(ctx, 3 * n);
    if (P.has_vel) P.vel.alloc(ctx, 3 * n);
    if (P.has_mass) P.mass.alloc(ctx, n);
    if (P.has_pid) P.pid.alloc(ctx, n);
    const int T = 256;
    UnpackArgs a{};
    a.stride = v.stride;
    a.off_pos = v.off_pos; a.off_vel = v.off_vel; a.off_mass = v.off_mass; a.off_pid = v.off_pid;
    a.pos = P.pos.p;
    a.vel = P.has_vel ? P.vel.p : nullptr;
    a.mass = P.has_mass ? P.mass.p : nullptr;
    a.pid = P.has_pid ? P.pid.p : nullptr;
    // Raw records are copied in chunks on a second stream into two device staging buffers (full PCIe rate for pinned host
    // memory; the driver bounces pageable memory itself) and unpacked on the context stream while the next chunk is in
    // flight.  Reading the caller's records zero-copy from the kernel instead measured 21 GB/s (profiles/r01a).
    if (!ctx->copy_stream) VR_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    KScope kup(ctx, "io.upload");
    // (A strided 2-D DMA of just the 64-byte window that holds the fields measured 2.5 GB/s against 54 GB/s for whole records.)
    const size_t width = v.stride;
    const size_t chunk_bytes = (size_t)64 << 20;
    size_t per = chunk_bytes / width;
    if (per < 1) per = 1;
    if (per > n) per = n;
    DBuf<unsigned char> stage[2];
    cudaEvent_t ready[2], freed[2];
    for (int b = 0; b < 2; b++) {
        stage[b].alloc(ctx, per * width);
        ready[b] = KScope::get(ctx);
        freed[b] = KScope::get(ctx);
    }
    // the staging buffers may still be in use by earlier work of the context stream
    VR_CUDA(cudaEventRecord(freed[0], ctx->stream));
    VR_CUDA(cudaEventRecord(freed[1], ctx->stream));
    int b = 0;
    for (size_t first = 0; first < n; first += per, b ^= 1) {
        const size_t cnt = n - first < per ? n - first : per;
        const unsigned char *src = static_cast<const unsigned char *>(v.base) + first * v.stride;
        VR_CUDA(cudaStreamWaitEvent(ctx->copy_stream, freed[b], 0));
        VR_CUDA(cudaMemcpyAsync(stage[b].p, src, cnt * v.stride, cudaMemcpyHostToDevice, ctx->copy_stream));
        VR_CUDA(cudaEventRecord(ready[b], ctx->copy_stream));
        VR_CUDA(cudaStreamWaitEvent(ctx->stream, ready[b], 0));
        a.base = stage[b].p;
        a.n = cnt;
        a.first = first;
        VR_LAUNCH(ctx, k_unpack, grid_for(cnt, T), T, 0, a);
        VR_CUDA(cudaEventRecord(freed[b], ctx->stream));
        ctx->stats.h2d_bytes += (long long)(cnt * width);
    }
    for (int k = 0; k < 2; k++) { ctx->evpool.push_back(ready[k]); ctx->evpool.push_back(freed[k]); }
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
