// velociraptor-stf_b200/csrc/api.cu -- the extern "C" boundary of libvrgpu.so (see include/vrgpu.h).
#include "fof.cuh"
#include "stages.cuh"

#include <chrono>
#include <cmath>
#include <cstdlib>

using namespace vrgpu;

namespace {

thread_local std::string g_init_error;

template <typename F> int guarded(vrgpu_ctx *ctx, F &&f)
{
    if (!ctx) return VRGPU_EINVAL;
    try {
        VR_CUDA(cudaSetDevice(ctx->device));
        f();
        return VRGPU_OK;
    } catch (const Error &e) {
        ctx->err = e.what();
        cudaGetLastError();
        return e.code;
    } catch (const std::bad_alloc &) {
        ctx->err = "host allocation failed";
        return VRGPU_ENOMEM;
    } catch (const std::exception &e) {
        ctx->err = e.what();
        return VRGPU_EINVAL;
    }
}

struct CallTimer {
    vrgpu_ctx *ctx;
    std::chrono::steady_clock::time_point t0;
    explicit CallTimer(vrgpu_ctx *c) : ctx(c), t0(std::chrono::steady_clock::now()) { c->stats = vrgpu_stats{}; }
    ~CallTimer()
    {
        cudaStreamSynchronize(ctx->stream);
        ctx->stats.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
};

// write a compact (index, xyz) list back into the caller's (writable) view
void scatter_positions_to_host(vrgpu_ctx *ctx, const vrgpu_view &view, size_t off_field, const uint32_t *idx_dev, const double *xyz_dev, size_t cnt)
{
    if (cnt == 0) return;
    std::vector<uint32_t> idx(cnt);
    std::vector<double> xyz(3 * cnt);
    VR_CUDA(cudaMemcpyAsync(idx.data(), idx_dev, cnt * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(xyz.data(), xyz_dev, 3 * cnt * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->stats.d2h_bytes += (long long)(cnt * 28);
    unsigned char *base = const_cast<unsigned char *>(static_cast<const unsigned char *>(view.base));
    for (size_t k = 0; k < cnt; k++) {
        double *p = reinterpret_cast<double *>(base + (size_t)idx[k] * view.stride + off_field);
        p[0] = xyz[3 * k]; p[1] = xyz[3 * k + 1]; p[2] = xyz[3 * k + 2];
    }
}

void upload_i64(vrgpu_ctx *ctx, const long long *host, size_t n, DBuf<long long> &dev)
{
    dev.alloc(ctx, n ? n : 1);
    if (n) VR_CUDA(cudaMemcpyAsync(dev.p, host, n * sizeof(long long), cudaMemcpyHostToDevice, ctx->stream));
    ctx->stats.h2d_bytes += (long long)(n * sizeof(long long));
}

} // namespace

extern "C" {

const char *vrgpu_version(void) { return "vrgpu 0.1 sm_100a"; }

int vrgpu_init(int device, vrgpu_ctx **out)
{
    if (!out) return VRGPU_EINVAL;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return VRGPU_ENODEV;
    }
    if (device < 0) {
        if (cudaGetDevice(&device) != cudaSuccess) device = 0;
    }
    if (device >= ndev) return VRGPU_EINVAL;
    vrgpu_ctx *ctx = new (std::nothrow) vrgpu_ctx;
    if (!ctx) return VRGPU_ENOMEM;
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { delete ctx; return VRGPU_ECUDA; }
    cudaDeviceProp prop{};
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return VRGPU_ECUDA; }
    if (prop.major < 10) {
        // the fat binary only holds sm_100a code
        delete ctx;
        return VRGPU_ENODEV;
    }
    ctx->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return VRGPU_ECUDA; }
    *out = ctx;
    return VRGPU_OK;
}

void vrgpu_destroy(vrgpu_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->pool.release_all();
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *vrgpu_last_error(const vrgpu_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

void vrgpu_get_stats(const vrgpu_ctx *ctx, vrgpu_stats *out)
{
    if (ctx && out) *out = ctx->stats;
}

int vrgpu_fof3d(vrgpu_ctx *ctx, vrgpu_view view, double b, const double *period, long long minsize, long long *pfof_out,
                long long *ngroups_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(pfof_out && ngroups_out, VRGPU_EINVAL, "null output pointer");
        if (period) VR_REQUIRE(period[0] == period[1] && period[1] == period[2], VRGPU_EUNSUPPORTED, "only cubic periods are supported");
        CallTimer ct(ctx);
        DeviceParticles P;
        Fof3dState S;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, false, false, true, P);
            ctx->stats.h2d_ms += t.stop();
        }
        fof3d_device(ctx, P, b, period ? period[0] : 0.0, minsize, S);
        {
            StageTimer t(ctx);
            download_i64(ctx, S.pfof.p, pfof_out, P.n);
            ctx->stats.d2h_ms += t.stop();
        }
        *ngroups_out = S.ngroups;
        ctx->stats.n_groups_final = S.ngroups;
    });
}

int vrgpu_adjust_for_period(vrgpu_ctx *ctx, vrgpu_view view, double period, const long long *pfof, long long ngroups)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(pfof != nullptr, VRGPU_EINVAL, "null pfof");
        CallTimer ct(ctx);
        DeviceParticles P;
        upload_view(ctx, view, false, false, false, P);
        DBuf<long long> pf;
        upload_i64(ctx, pfof, P.n, pf);
        DBuf<uint32_t> midx;
        DBuf<double> mxyz;
        size_t cnt = adjust_for_period_device(ctx, P, pf.p, ngroups, period, midx, mxyz);
        scatter_positions_to_host(ctx, view, view.off_pos, midx.p, mxyz.p, cnt);
    });
}

int vrgpu_fof6d(vrgpu_ctx *ctx, vrgpu_view view, const long long *pfof3d, long long ngroups3d, double b6, double vfac, int mode,
                long long minsize, long long *pfof_out, long long *ngroups_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(pfof3d && pfof_out && ngroups_out, VRGPU_EINVAL, "null pointer");
        VR_REQUIRE(mode == 3 || mode == 4, VRGPU_EINVAL, "mode must be 3 (FOF6DADAPTIVE) or 4 (FOF6D)");
        CallTimer ct(ctx);
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, true, true, true, P);
            ctx->stats.h2d_ms += t.stop();
        }
        DBuf<long long> pf3, pf6;
        upload_i64(ctx, pfof3d, P.n, pf3);
        pf6.alloc(ctx, P.n ? P.n : 1);
        long long ng = 0;
        {
            StageTimer t(ctx);
            ng = fof6d_device(ctx, P, pf3.p, ngroups3d, b6, vfac, mode, minsize, pf6.p, true);
            ctx->stats.fof6d_ms += t.stop();
        }
        {
            StageTimer t(ctx);
            writeback_phase(ctx, view, P, pf3.p);
            download_i64(ctx, pf6.p, pfof_out, P.n);
            ctx->stats.d2h_ms += t.stop();
        }
        *ngroups_out = ng;
        ctx->stats.n_groups_final = ng;
    });
}

int vrgpu_potential(vrgpu_ctx *ctx, vrgpu_view view, long long ngroups, const long long *offsets, const long long *index, double G,
                    double eps, double theta, int bucket, long long pp_threshold, double *phi_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(ngroups >= 0 && offsets && phi_out, VRGPU_EINVAL, "bad arguments");
        CallTimer ct(ctx);
        if (ngroups == 0) return;
        const long long ntot = offsets[ngroups];
        VR_REQUIRE(ntot >= 0 && (ntot == 0 || index), VRGPU_EINVAL, "bad CSR");
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, false, true, false, P);
            ctx->stats.h2d_ms += t.stop();
        }
        potential_csr_host(ctx, P, ngroups, offsets, index, G, eps, theta, bucket, pp_threshold, phi_out);
    });
}

int vrgpu_unbind(vrgpu_ctx *ctx, vrgpu_view view, long long *pfof_inout, long long *ngroups_inout, const vrgpu_unbind_opts *opts,
                 int *changed_out, long long *n_marginal_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(pfof_inout && ngroups_inout && opts, VRGPU_EINVAL, "null pointer");
        CallTimer ct(ctx);
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, true, true, true, P);
            ctx->stats.h2d_ms += t.stop();
        }
        DBuf<long long> pf;
        upload_i64(ctx, pfof_inout, P.n, pf);
        long long ng = *ngroups_inout;
        int changed = 0;
        long long nmarg = 0;
        unbind_device(ctx, P, pf.p, ng, *opts, changed, nmarg);
        {
            StageTimer t(ctx);
            download_i64(ctx, pf.p, pfof_inout, P.n);
            ctx->stats.d2h_ms += t.stop();
        }
        *ngroups_inout = ng;
        if (changed_out) *changed_out = changed;
        if (n_marginal_out) *n_marginal_out = nmarg;
        ctx->stats.n_groups_final = ng;
        ctx->stats.n_marginal = nmarg;
    });
}

int vrgpu_search_full_set(vrgpu_ctx *ctx, vrgpu_view view, const vrgpu_search_opts *o, long long *pfof_out, long long *ngroups_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(o && pfof_out && ngroups_out, VRGPU_EINVAL, "null pointer");
        VR_REQUIRE(o->fofbgtype == 5 || o->fofbgtype == 4 || o->fofbgtype == 3, VRGPU_EUNSUPPORTED,
                   "FoF_Field_search_type must be 5 (3DFOF), 4 (6DFOF) or 3 (adaptive 6DFOF)");
        CallTimer ct(ctx);
        const bool six = o->fofbgtype <= 4;
        const bool unbind = o->bound_halos >= 1;
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, six || unbind, six || unbind, true, P);
            ctx->stats.h2d_ms += t.stop();
        }
        Fof3dState S;
        fof3d_device(ctx, P, std::sqrt(o->param1), o->period, o->halo_min_size, S);
        long long ng = S.ngroups;
        DBuf<long long> pf = std::move(S.pfof);
        // search.cxx:393-396
        if (o->period > 0 && ng > 0) {
            DBuf<uint32_t> midx;
            DBuf<double> mxyz;
            size_t cnt = adjust_for_period_device(ctx, P, pf.p, ng, o->period, midx, mxyz);
            StageTimer t(ctx);
            scatter_positions_to_host(ctx, view, view.off_pos, midx.p, mxyz.p, cnt);
            ctx->stats.d2h_ms += t.stop();
        }
        // free the grid before the later stages need memory
        S = Fof3dState{};
        if (six && ng > 0) {
            DBuf<long long> pf6(ctx, P.n);
            StageTimer t(ctx);
            // search.cxx:410: param[1] *= ellhalo6dxfac^2 ; b6 = sqrt(param[1])
            double param1 = o->param1 * (o->ellhalo6dxfac * o->ellhalo6dxfac);
            long long ng6 = fof6d_device(ctx, P, pf.p, ng, std::sqrt(param1), o->ellhalo6dvfac, o->fofbgtype, o->halo_min_size, pf6.p, true);
            ctx->stats.fof6d_ms += t.stop();
            {
                StageTimer t2(ctx);
                writeback_phase(ctx, view, P, pf.p);
                ctx->stats.d2h_ms += t2.stop();
            }
            pf = std::move(pf6);
            ng = ng6;
        }
        if (unbind && ng > 0) {
            int changed = 0;
            long long nmarg = 0;
            unbind_device(ctx, P, pf.p, ng, o->unbind, changed, nmarg);
            ctx->stats.n_marginal = nmarg;
        }
        {
            StageTimer t(ctx);
            download_i64(ctx, pf.p, pfof_out, P.n);
            ctx->stats.d2h_ms += t.stop();
        }
        *ngroups_out = ng;
        ctx->stats.n_groups_final = ng;
    });
}

} // extern "C"
