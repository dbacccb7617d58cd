// velociraptor-stf_b200/csrc/io.cu -- host <-> device marshalling of the caller's strided particle view.
//
// The reference holds particles as a 96-byte AoS (NBody::Particle, NBodylib/src/NBody/Particle.h:75-109); the
// device wants packed/SoA columns.  Two paths, chosen per call:
//   * pinned / registered host memory  -> the unpack kernel reads the caller's records directly over PCIe
//                                          (zero-copy): only the sectors holding the needed fields move.
//   * pageable host memory             -> chunks of the raw records are copied with cudaMemcpyAsync into a
//                                          device staging buffer (the driver stages through its own pinned
//                                          buffers) and unpacked there.
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

static bool host_ptr_is_pinned(const void *p, void **devptr)
{
    cudaPointerAttributes at{};
    cudaError_t e = cudaPointerGetAttributes(&at, p);
    if (e != cudaSuccess) { cudaGetLastError(); return false; }
    if (at.type == cudaMemoryTypeHost && at.devicePointer) { *devptr = at.devicePointer; return true; }
    return false;
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
    // bytes of each record that this call actually needs (32-byte sector granularity for the zero-copy path)
    size_t fields = 24 + (P.has_vel ? 24 : 0) + (P.has_mass ? 8 : 0) + (P.has_pid ? 8 : 0);
    void *dev = nullptr;
    if (host_ptr_is_pinned(v.base, &dev)) {
        a.base = static_cast<const unsigned char *>(dev);
        a.n = n;
        a.first = 0;
        VR_LAUNCH(ctx, k_unpack, grid_for(n, T), T, 0, a);
        ctx->stats.h2d_bytes += (long long)(n * fields);
    } else {
        // pageable: stage raw records through the device in chunks
        const size_t chunk_bytes = (size_t)256 << 20;
        size_t per = chunk_bytes / v.stride;
        if (per < 1) per = 1;
        if (per > n) per = n;
        DBuf<unsigned char> stage(ctx, per * v.stride);
        for (size_t first = 0; first < n; first += per) {
            size_t cnt = n - first < per ? n - first : per;
            VR_CUDA(cudaMemcpyAsync(stage.p, static_cast<const unsigned char *>(v.base) + first * v.stride, cnt * v.stride,
                                    cudaMemcpyHostToDevice, ctx->stream));
            a.base = stage.p;
            a.n = cnt;
            a.first = first;
            VR_LAUNCH(ctx, k_unpack, grid_for(cnt, T), T, 0, a);
            ctx->stats.h2d_bytes += (long long)(cnt * v.stride);
        }
    }
}

void download_i64(vrgpu_ctx *ctx, const long long *dev, long long *host, size_t n)
{
    if (n == 0) return;
    VR_CUDA(cudaMemcpyAsync(host, dev, n * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->stats.d2h_bytes += (long long)(n * sizeof(long long));
}

} // namespace vrgpu
