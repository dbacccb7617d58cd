// velociraptor-stf_b200/csrc/io.cu -- host <-> device marshalling of the caller's strided particle view.
//
// The reference holds particles as a 96-byte AoS (NBody::Particle, NBodylib/src/NBody/Particle.h:75-109); the
// device wants packed columns.  Chunks of the raw records are copied with cudaMemcpyAsync (copy stream, double
// buffered) into device staging buffers and unpacked there by k_unpack on the context stream.
#include "fof.cuh"

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
    const size_t chunk_bytes = (size_t)64 << 20;
    size_t per = chunk_bytes / v.stride;
    if (per < 1) per = 1;
    if (per > n) per = n;
    DBuf<unsigned char> stage[2];
    cudaEvent_t ready[2], freed[2];
    for (int b = 0; b < 2; b++) {
        stage[b].alloc(ctx, per * v.stride);
        ready[b] = KScope::get(ctx);
        freed[b] = KScope::get(ctx);
    }
    // the staging buffers may still be in use by earlier work of the context stream
    VR_CUDA(cudaEventRecord(freed[0], ctx->stream));
    VR_CUDA(cudaEventRecord(freed[1], ctx->stream));
    int b = 0;
    for (size_t first = 0; first < n; first += per, b ^= 1) {
        const size_t cnt = n - first < per ? n - first : per;
        VR_CUDA(cudaStreamWaitEvent(ctx->copy_stream, freed[b], 0));
        VR_CUDA(cudaMemcpyAsync(stage[b].p, static_cast<const unsigned char *>(v.base) + first * v.stride, cnt * v.stride,
                                cudaMemcpyHostToDevice, ctx->copy_stream));
        VR_CUDA(cudaEventRecord(ready[b], ctx->copy_stream));
        VR_CUDA(cudaStreamWaitEvent(ctx->stream, ready[b], 0));
        a.base = stage[b].p;
        a.n = cnt;
        a.first = first;
        VR_LAUNCH(ctx, k_unpack, grid_for(cnt, T), T, 0, a);
        VR_CUDA(cudaEventRecord(freed[b], ctx->stream));
        ctx->stats.h2d_bytes += (long long)(cnt * v.stride);
    }
    for (int k = 0; k < 2; k++) { ctx->evpool.push_back(ready[k]); ctx->evpool.push_back(freed[k]); }
}

void download_i64(vrgpu_ctx *ctx, const long long *dev, long long *host, size_t n)
{
    if (n == 0) return;
    VR_CUDA(cudaMemcpyAsync(host, dev, n * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->stats.d2h_bytes += (long long)(n * sizeof(long long));
}

} // namespace vrgpu
