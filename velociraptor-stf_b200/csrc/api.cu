// velociraptor-stf_b200/csrc/api.cu -- the extern "C" boundary of libvrgpu.so (see include/vrgpu.h).
#include "fof.cuh"
#include "groups.cuh"
#include "stages.cuh"

#include <chrono>
#include <cmath>
#include <cstdlib>
#include <memory>

using namespace vrgpu;

namespace vrgpu {
struct ResidentState {
    DeviceParticles pristine;
    DBuf<long long> pfof;
    long long ngroups = 0;
};
} // namespace vrgpu

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
        resolve_ktimes(ctx);
        ctx->stats.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
};

// write a compact (index, xyz) list back into the caller's (writable) view
__global__ void k_scatter_direct(unsigned char *__restrict__ base, size_t stride, size_t off_field, const uint32_t *__restrict__ idx,
                                 const double *__restrict__ xyz, size_t cnt)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= cnt) return;
    double *p = reinterpret_cast<double *>(base + (size_t)idx[k] * stride + off_field);
    p[0] = xyz[3 * k]; p[1] = xyz[3 * k + 1]; p[2] = xyz[3 * k + 2];
}

void scatter_positions_to_host(vrgpu_ctx *ctx, const vrgpu_view &view, size_t off_field, const uint32_t *idx_dev, const double *xyz_dev, size_t cnt)
{
    if (cnt == 0) return;
    if (void *dev = mapped_host_pointer(view.base)) {
        // page-locked caller memory: the device writes the moved positions straight into the records
        VR_LAUNCH(ctx, k_scatter_direct, grid_for(cnt, 256), 256, 0, static_cast<unsigned char *>(dev), view.stride, off_field, idx_dev, xyz_dev, cnt);
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->stats.d2h_bytes += (long long)(cnt * 24);
        return;
    }
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


void check_search_opts(const vrgpu_search_opts *o)
{
    VR_REQUIRE(o->fofbgtype == 5 || o->fofbgtype == 4 || o->fofbgtype == 3, VRGPU_EUNSUPPORTED,
               "FoF_Field_search_type must be 5 (3DFOF), 4 (6DFOF) or 3 (adaptive 6DFOF)");
    VR_REQUIRE(o->param1 > 0, VRGPU_EINVAL, "param1 (squared linking length) must be positive");
}

// Borrow read-only columns (vel / mass / pid) of the pristine resident copy into a working DeviceParticles without
// copying; the borrowed DBufs are detached again before they could be returned to the pool twice.
struct BorrowGuard {
    DeviceParticles &P;
    bool vel;
    BorrowGuard(DeviceParticles &p, const DeviceParticles &S, bool borrow_vel) : P(p), vel(borrow_vel && S.has_vel)
    {
        if (vel) { P.vel.p = S.vel.p; P.vel.n = S.vel.n; P.vel.ctx = nullptr; P.has_vel = true; }
        if (S.has_mass) { P.mass.p = S.mass.p; P.mass.n = S.mass.n; P.mass.ctx = nullptr; P.has_mass = true; }
        if (S.has_pid) { P.pid.p = S.pid.p; P.pid.n = S.pid.n; P.pid.ctx = nullptr; P.has_pid = true; }
    }
    ~BorrowGuard()
    {
        if (vel) { P.vel.p = nullptr; P.vel.n = 0; }
        if (P.mass.ctx == nullptr) { P.mass.p = nullptr; P.mass.n = 0; }
        if (P.pid.ctx == nullptr) { P.pid.p = nullptr; P.pid.n = 0; }
    }
};

// The DM path of SearchFullSet (src/search.cxx:24-854) on device-resident particles.  `view` (may be null) is the
// caller's writable host array that receives the reference's in-place coordinate updates.
void run_search(vrgpu_ctx *ctx, DeviceParticles &P, const vrgpu_search_opts *o, const vrgpu_view *view, DBuf<long long> &pf, long long &ng)
{
    const bool six = o->fofbgtype <= 4;
    const bool unbind = o->bound_halos >= 1;
    // the asynchronous write-back (experiment, see below) reads pf3_keep / P until `ev`: joined on every way out
    DBuf<long long> pf3_keep;
    struct WbJoin {
        vrgpu_ctx *ctx;
        cudaEvent_t ev = nullptr;
        HostWriteback *host = nullptr; // pageable caller memory: a worker thread scatters the coordinates beside the unbinding stage
        ~WbJoin()
        {
            hostwb_join(ctx, host, false);
            if (!ev) return;
            cudaStreamWaitEvent(ctx->stream, ev, 0);
            cudaEventSynchronize(ev);
            ctx->evpool.push_back(ev);
        }
    } wb{ctx};
    Fof3dState S;
    fof3d_device(ctx, P, std::sqrt(o->param1), o->period, o->halo_min_size, S);
    ng = S.ngroups;
    pf = std::move(S.pfof);
    // search.cxx:393-396
    if (o->period > 0 && ng > 0) {
        DBuf<uint32_t> midx;
        DBuf<double> mxyz;
        size_t cnt;
        {
            KScope k(ctx, "adjust_for_period");
            cnt = adjust_for_period_device(ctx, P, pf.p, ng, o->period, midx, mxyz);
        }
        if (view) {
            StageTimer t(ctx);
            scatter_positions_to_host(ctx, *view, view->off_pos, midx.p, mxyz.p, cnt);
            ctx->stats.d2h_ms += t.stop();
        }
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
        if (view) {
            // experiment (off by default): the write-back runs on the copy stream beside the unbinding stage
            static const bool async_wb = [] { const char *e = getenv("VRGPU_ASYNC_WRITEBACK"); return e && atoi(e) == 1; }();
            // pageable records (what the reference passes): on by default when an unbinding stage follows, VRGPU_ASYNC_WRITEBACK=0 keeps
            // it in front of that stage (profiles/r02_e2e.md)
            static const bool host_async = [] { const char *e = getenv("VRGPU_ASYNC_WRITEBACK"); return !(e && atoi(e) == 0); }();
            if (async_wb) wb.ev = writeback_phase_async(ctx, *view, P, pf.p);
            if (!wb.ev && host_async && unbind) wb.host = hostwb_start(ctx, *view, P, pf.p);
            if (wb.ev) pf3_keep = std::move(pf); // read by the copy stream until wb.ev
            else if (wb.host) {}
            else {
                StageTimer t2(ctx);
                writeback_phase(ctx, *view, P, pf.p);
                ctx->stats.d2h_ms += t2.stop();
            }
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
    if (wb.host) { // errors of the worker surface here; on an exception above the destructor joins it silently
        StageTimer t2(ctx);
        HostWriteback *h = wb.host;
        wb.host = nullptr;
        hostwb_join(ctx, h, true);
        ctx->stats.d2h_ms += t2.stop();
    }
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
    delete ctx->resident;
    ctx->resident = nullptr;
    mg_destroy(ctx);
    resolve_ktimes(ctx);
    for (cudaEvent_t e : ctx->evpool) cudaEventDestroy(e);
    ctx->evpool.clear();
    ctx->pool.release_all();
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    for (int k = 0; k < 4; k++) if (ctx->aux_stream[k]) cudaStreamDestroy(ctx->aux_stream[k]);
    for (int k = 0; k < 2; k++) if (ctx->stage_host[k]) cudaFreeHost(ctx->stage_host[k]);
    for (int k = 0; k < 3; k++) if (ctx->up_host[k]) cudaFreeHost(ctx->up_host[k]);
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

int vrgpu_fof_criterion6d(vrgpu_ctx *ctx, vrgpu_view view, const double *check, double check_below, double xlink2, double vlink2,
                          long long minsize, long long *pfof_out, long long *ngroups_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(pfof_out && ngroups_out, VRGPU_EINVAL, "null pointer");
        CallTimer ct(ctx);
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, true, false, true, P);
            ctx->stats.h2d_ms += t.stop();
        }
        DBuf<double> chk;
        if (check && P.n) {
            chk.alloc(ctx, P.n);
            VR_CUDA(cudaMemcpyAsync(chk.p, check, P.n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
            ctx->stats.h2d_bytes += (long long)(P.n * sizeof(double));
        }
        DBuf<long long> pf(ctx, P.n ? P.n : 1);
        long long ng = 0;
        {
            StageTimer t(ctx);
            ng = fof_criterion6d_device(ctx, P, check ? chk.p : nullptr, check_below, xlink2, vlink2, minsize, pf.p);
            ctx->stats.fof6d_ms += t.stop();
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

int vrgpu_velocity_density(vrgpu_ctx *ctx, vrgpu_view view, const int *type, const long long *leaf_offsets, long long nleaves, double period,
                           int bsize, int nsearch, int nvel, int approx, double *density_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(density_out, VRGPU_EINVAL, "null pointer");
        if (leaf_offsets) {
            VR_REQUIRE(nleaves > 0 && nleaves < 0x7fffffffll && leaf_offsets[0] == 0 && leaf_offsets[nleaves] == (long long)view.n, VRGPU_EINVAL,
                       "leaf offsets must run from 0 to the number of particles");
            for (long long l = 0; l < nleaves; l++)
                VR_REQUIRE(leaf_offsets[l] <= leaf_offsets[l + 1], VRGPU_EINVAL, "leaf offsets must not decrease");
        }
        CallTimer ct(ctx);
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, true, false, false, P);
            ctx->stats.h2d_ms += t.stop();
        }
        if (P.n == 0) return;
        DBuf<int> tp;
        if (type) {
            tp.alloc(ctx, P.n);
            VR_CUDA(cudaMemcpyAsync(tp.p, type, P.n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
            ctx->stats.h2d_bytes += (long long)(P.n * sizeof(int));
        }
        DBuf<long long> loff;
        if (leaf_offsets) {
            loff.alloc(ctx, (size_t)nleaves + 1);
            VR_CUDA(cudaMemcpyAsync(loff.p, leaf_offsets, ((size_t)nleaves + 1) * sizeof(long long), cudaMemcpyHostToDevice, ctx->stream));
            ctx->stats.h2d_bytes += (long long)(((size_t)nleaves + 1) * sizeof(long long));
        }
        DBuf<double> den(ctx, P.n);
        velocity_density_device(ctx, P, type ? tp.p : nullptr, leaf_offsets ? loff.p : nullptr, leaf_offsets ? nleaves : 0, period, bsize, nsearch,
                                nvel, approx, den.p);
        {
            StageTimer t(ctx);
            VR_CUDA(cudaMemcpyAsync(density_out, den.p, P.n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
            ctx->stats.d2h_bytes += (long long)(P.n * sizeof(double));
            ctx->stats.d2h_ms += t.stop();
        }
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

int vrgpu_tree_order(vrgpu_ctx *ctx, vrgpu_view view, long long ngroups, const long long *offsets, const long long *index, int bucket,
                     long long *perm_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(ngroups >= 0 && offsets && perm_out, VRGPU_EINVAL, "bad arguments");
        CallTimer ct(ctx);
        if (ngroups == 0) return;
        VR_REQUIRE(offsets[ngroups] >= 0 && (offsets[ngroups] == 0 || index), VRGPU_EINVAL, "bad CSR");
        for (long long g = 0; g < ngroups; g++) VR_REQUIRE(offsets[g + 1] > offsets[g], VRGPU_EINVAL, "empty group");
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, false, false, false, P);
            ctx->stats.h2d_ms += t.stop();
        }
        tree_order_csr_host(ctx, P, ngroups, offsets, index, bucket, perm_out);
    });
}

int vrgpu_binding_energy(vrgpu_ctx *ctx, vrgpu_view view, long long ngroups, const long long *offsets, const double *gcmvel, double G,
                         double eps, double theta, int bucket, long long large_threshold, double *phi_out, double *E_out, double *T_out,
                         double *Pot_out, double *Efrac_out, long long *iminpot_out, long long *ibound_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(ngroups >= 0 && offsets, VRGPU_EINVAL, "bad arguments");
        CallTimer ct(ctx);
        if (ngroups == 0) return;
        VR_REQUIRE(gcmvel && T_out && Pot_out && Efrac_out && iminpot_out && ibound_out, VRGPU_EINVAL, "null per-group array");
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, true, true, false, P);
            ctx->stats.h2d_ms += t.stop();
        }
        binding_energy_host(ctx, P, ngroups, offsets, gcmvel, G, eps, theta, bucket, large_threshold, phi_out, E_out, T_out, Pot_out,
                            Efrac_out, iminpot_out, ibound_out);
    });
}

int vrgpu_sort_groups_by_key(vrgpu_ctx *ctx, long long ngroups, const long long *offsets, const double *key, long long *order_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(ngroups >= 0 && offsets, VRGPU_EINVAL, "bad arguments");
        CallTimer ct(ctx);
        if (ngroups == 0 || offsets[ngroups] == 0) return;
        VR_REQUIRE(key && order_out, VRGPU_EINVAL, "null key / order array");
        sort_groups_host(ctx, ngroups, offsets, key, order_out);
    });
}

int vrgpu_build_pglist(vrgpu_ctx *ctx, const long long *pfof, long long n, long long ngroups, long long *numingroup_out,
                       long long *index_out, long long *ngrouped_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(numingroup_out && index_out && ngrouped_out, VRGPU_EINVAL, "null output pointer");
        CallTimer ct(ctx);
        if (pfof == nullptr) {
            VR_REQUIRE(ctx->resident && ctx->resident->pfof.p, VRGPU_EINVAL, "no resident result");
            *ngrouped_out = pglist_device(ctx, ctx->resident->pfof.p, ctx->resident->pristine.n, ctx->resident->ngroups, numingroup_out, index_out);
            return;
        }
        VR_REQUIRE(n >= 0 && ngroups >= 0, VRGPU_EINVAL, "negative count");
        DBuf<long long> pf;
        {
            StageTimer t(ctx);
            upload_i64(ctx, pfof, (size_t)n, pf);
            ctx->stats.h2d_ms += t.stop();
        }
        *ngrouped_out = pglist_device(ctx, pf.p, (size_t)n, ngroups, numingroup_out, index_out);
        if (ngroups > 0 && n == 0) for (long long g = 0; g <= ngroups; g++) numingroup_out[g] = 0;
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
        check_search_opts(o);
        CallTimer ct(ctx);
        const bool six = o->fofbgtype <= 4;
        const bool unbind = o->bound_halos >= 1;
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, six || unbind, six || unbind, true, P);
            ctx->stats.h2d_ms += t.stop();
        }
        DBuf<long long> pf;
        long long ng = 0;
        run_search(ctx, P, o, &view, pf, ng);
        {
            StageTimer t(ctx);
            download_i64(ctx, pf.p, pfof_out, P.n);
            ctx->stats.d2h_ms += t.stop();
        }
        *ngroups_out = ng;
        ctx->stats.n_groups_final = ng;
    });
}

int vrgpu_particles_upload(vrgpu_ctx *ctx, vrgpu_view view)
{
    return guarded(ctx, [&] {
        CallTimer ct(ctx);
        delete ctx->resident;
        ctx->resident = nullptr;
        std::unique_ptr<ResidentState> R(new ResidentState);
        StageTimer t(ctx);
        upload_view(ctx, view, view.off_vel != VRGPU_NOFIELD, view.off_mass != VRGPU_NOFIELD, view.off_pid != VRGPU_NOFIELD, R->pristine);
        ctx->stats.h2d_ms += t.stop();
        ctx->resident = R.release();
    });
}

int vrgpu_particles_release(vrgpu_ctx *ctx)
{
    return guarded(ctx, [&] {
        delete ctx->resident;
        ctx->resident = nullptr;
        ctx->pool.release_cached();
    });
}

int vrgpu_search_full_set_resident(vrgpu_ctx *ctx, const vrgpu_search_opts *o, long long *ngroups_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(o && ngroups_out, VRGPU_EINVAL, "null pointer");
        VR_REQUIRE(ctx->resident != nullptr, VRGPU_EINVAL, "no resident particles: call vrgpu_particles_upload first");
        check_search_opts(o);
        CallTimer ct(ctx);
        ResidentState &R = *ctx->resident;
        const DeviceParticles &S = R.pristine;
        const bool six = o->fofbgtype <= 4;
        const bool unbind = o->bound_halos >= 1;
        VR_REQUIRE(!(six || unbind) || S.has_vel, VRGPU_EINVAL, "resident particles were uploaded without velocities");
        // working copy: the path rewrites positions (AdjustStructureForPeriod) and phase-space coordinates (6DFOF)
        DeviceParticles P;
        P.n = S.n;
        {
            KScope k(ctx, "copy_pristine");
            P.pos.alloc(ctx, 3 * S.n + 1);
            VR_CUDA(cudaMemcpyAsync(P.pos.p, S.pos.p, 3 * S.n * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
            if (six && S.has_vel) {
                P.vel.alloc(ctx, 3 * S.n + 1);
                VR_CUDA(cudaMemcpyAsync(P.vel.p, S.vel.p, 3 * S.n * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
                P.has_vel = true;
            }
        }
        // columns the path only reads are borrowed from the pristine copy
        BorrowGuard bg(P, S, !(six && S.has_vel) && (six || unbind));
        run_search(ctx, P, o, nullptr, R.pfof, R.ngroups);
        *ngroups_out = R.ngroups;
        ctx->stats.n_groups_final = R.ngroups;
    });
}

int vrgpu_resident_pfof(vrgpu_ctx *ctx, long long *pfof_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(pfof_out, VRGPU_EINVAL, "null pointer");
        VR_REQUIRE(ctx->resident && ctx->resident->pfof.p, VRGPU_EINVAL, "no resident result");
        download_i64(ctx, ctx->resident->pfof.p, pfof_out, ctx->resident->pristine.n);
    });
}

int vrgpu_dist_unique_id(char *id_out)
{
    if (!id_out) return VRGPU_EINVAL;
    try {
        mg_unique_id(id_out);
        return VRGPU_OK;
    } catch (const Error &e) {
        g_init_error = e.what();
        return e.code;
    }
}

int vrgpu_dist_init(vrgpu_ctx *ctx, int nranks, int rank, const char *id)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(id != nullptr || nranks == 1, VRGPU_EINVAL, "null unique id");
        mg_init(ctx, nranks, rank, id);
    });
}

int vrgpu_dist_search_full_set(vrgpu_ctx *ctx, vrgpu_view view, double slab_lo, double slab_hi, const vrgpu_search_opts *o,
                               long long *pfof_out, long long *ngroups_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(o && pfof_out && ngroups_out, VRGPU_EINVAL, "null pointer");
        check_search_opts(o);
        CallTimer ct(ctx);
        const bool six = o->fofbgtype <= 4;
        const bool unbind = o->bound_halos >= 1;
        DeviceParticles P;
        {
            StageTimer t(ctx);
            upload_view(ctx, view, six || unbind, six || unbind, true, P);
            ctx->stats.h2d_ms += t.stop();
        }
        DBuf<long long> pf;
        long long ng = 0;
        // the view receives the reference's in-place coordinate updates, as in vrgpu_search_full_set
        if (mg_nranks(ctx) == 1) run_search(ctx, P, o, &view, pf, ng);
        else {
            MgWriteback wb;
            mg_search(ctx, P, slab_lo, slab_hi, o, pf, ng, &wb);
            StageTimer t(ctx);
            if (wb.phase) writeback_phase(ctx, view, P, wb.pf3.p);
            else if (wb.nmoved) scatter_positions_to_host(ctx, view, view.off_pos, wb.midx.p, wb.mxyz.p, wb.nmoved);
            ctx->stats.d2h_ms += t.stop();
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

int vrgpu_dist_search_full_set_resident(vrgpu_ctx *ctx, double slab_lo, double slab_hi, const vrgpu_search_opts *o, long long *ngroups_out)
{
    if (ctx && mg_nranks(ctx) == 1) return vrgpu_search_full_set_resident(ctx, o, ngroups_out);
    return guarded(ctx, [&] {
        VR_REQUIRE(o && ngroups_out, VRGPU_EINVAL, "null pointer");
        VR_REQUIRE(ctx->resident != nullptr, VRGPU_EINVAL, "no resident particles: call vrgpu_particles_upload first");
        check_search_opts(o);
        CallTimer ct(ctx);
        ResidentState &R = *ctx->resident;
        // the multi-GPU path never writes into its input columns (it works on the extended / consolidated copies)
        mg_search(ctx, R.pristine, slab_lo, slab_hi, o, R.pfof, R.ngroups);
        *ngroups_out = R.ngroups;
        ctx->stats.n_groups_final = R.ngroups;
    });
}

void *vrgpu_stream(vrgpu_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

const char *vrgpu_kernel_times(vrgpu_ctx *ctx)
{
    if (!ctx) return "";
    resolve_ktimes(ctx);
    ctx->ktimes_text.clear();
    char line[256];
    for (auto &kv : ctx->ktimes) {
        snprintf(line, sizeof line, "%s %.6f %lld\n", kv.first.c_str(), kv.second.ms, kv.second.count);
        ctx->ktimes_text += line;
    }
    return ctx->ktimes_text.c_str();
}

int vrgpu_measure_fp64_tflops(vrgpu_ctx *ctx, double *tflops_out)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(tflops_out, VRGPU_EINVAL, "null pointer");
        *tflops_out = measure_fp64_tflops(ctx);
    });
}

int vrgpu_probe_fp64_pipes(vrgpu_ctx *ctx, double *out4)
{
    return guarded(ctx, [&] {
        VR_REQUIRE(out4, VRGPU_EINVAL, "null pointer");
        probe_fp64_pipes(ctx, out4);
    });
}

void vrgpu_kernel_times_reset(vrgpu_ctx *ctx)
{
    if (!ctx) return;
    resolve_ktimes(ctx);
    ctx->ktimes.clear();
}

} // extern "C"
