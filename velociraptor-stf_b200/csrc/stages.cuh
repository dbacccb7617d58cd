// velociraptor-stf_b200/csrc/stages.cuh -- stage entry points behind the C ABI (6DFOF, potentials, unbinding).
#pragma once
#include "fof.cuh"

namespace vrgpu {

// 6DFOF inside each 3D group (src/search.cxx:400-698).  pfof3d_dev / pfof6d_dev are device arrays in ORIGINAL
// particle order.  When apply_q3 is set, P.pos / P.vel receive the reference's scale-unscale rounding.
long long fof6d_device(vrgpu_ctx *ctx, DeviceParticles &P, const long long *pfof3d_dev, long long ngroups3d, double b6, double vfac,
                       int mode, long long minsize, long long *pfof6d_dev, bool apply_q3);

// write pos+vel of every particle with pfof_dev[i] > 0 back into the caller's writable view
long long fof_criterion6d_device(vrgpu_ctx *ctx, const DeviceParticles &P, const double *check_dev, double check_below, double p6, double p7,
                                 long long minsize, long long *pfof_dev);
void writeback_phase(vrgpu_ctx *ctx, const vrgpu_view &view, const DeviceParticles &P, const long long *pfof_dev);
cudaEvent_t writeback_phase_async(vrgpu_ctx *ctx, const vrgpu_view &view, const DeviceParticles &P, const long long *pfof_dev);
// The same into PAGEABLE caller memory, beside the stages that follow: the coordinates are packed on the context's stream (so P
// may change afterwards), a worker thread brings the packed records back on the copy stream through the page-locked staging
// buffers and scatters them into the caller's records while the caller's thread goes on launching the unbinding stage.
// hostwb_start returns false (nothing started) when the caller's memory is page-locked or nothing is grouped; hostwb_join must
// run before anything else uses the staging buffers and before the call returns.
struct HostWriteback;
HostWriteback *hostwb_start(vrgpu_ctx *ctx, const vrgpu_view &view, const DeviceParticles &P, const long long *pfof_dev);
void hostwb_join(vrgpu_ctx *ctx, HostWriteback *job, bool may_throw);

// GetVelocityDensity (src/localfield.cxx:18) for one structure; type_dev (original order) may be null; den_dev in original order (veldens.cu)
// leaf_off_dev (ncaller_leaves + 1 offsets into the caller's order, may be null): leaves of the caller's own tree for the per-leaf search
void velocity_density_device(vrgpu_ctx *ctx, const DeviceParticles &P, const int *type_dev, const long long *leaf_off_dev, long long ncaller_leaves,
                             double period, int bsize, int nsearch, int nvel, int approx, double *den_dev);

// device-visible alias of a page-locked host pointer, or nullptr for pageable memory (fof6d.cu)
void *mapped_host_pointer(const void *p);

// DFMA throughput of the device in TFLOP/s (potential.cu)
double measure_fp64_tflops(vrgpu_ctx *ctx);
void probe_fp64_pipes(vrgpu_ctx *ctx, double out4[4]);

// potentials for host-side CSR groups (vrgpu_potential)
void potential_csr_host(vrgpu_ctx *ctx, const DeviceParticles &P, long long ngroups, const long long *offsets_host,
                        const long long *index_host, double G, double eps, double theta, int bucket, long long pp_threshold,
                        double *phi_out_host);

// GetBindingEnergy (src/substructureproperties.cxx:3884-4255) for group-ordered particles (energy.cu)
void tree_order_csr_host(vrgpu_ctx *ctx, const DeviceParticles &P, long long ngroups, const long long *offsets_host,
                         const long long *index_host, int bucket, long long *perm_out_host);
void binding_energy_host(vrgpu_ctx *ctx, const DeviceParticles &P, long long ngroups, const long long *offsets_host,
                         const double *gcmvel_host, double G, double eps, double theta, int bucket, long long large_threshold,
                         double *phi_out, double *E_out, double *T_out, double *Pot_out, double *Efrac_out, long long *iminpot_out,
                         long long *ibound_out);

// per-group ascending order of a host key array (the qsort of SortAccordingtoBindingEnergy, substructureproperties.cxx:4302-4310)
void sort_groups_host(vrgpu_ctx *ctx, long long ngroups, const long long *offsets_host, const double *key_host, long long *order_out);

// CheckUnboundGroups + Unbind (src/unbind.cxx:196-304, 320-729) on device-resident particles
void unbind_device(vrgpu_ctx *ctx, const DeviceParticles &P, long long *pfof_dev, long long &ngroups, const vrgpu_unbind_opts &o,
                   int &changed, long long &n_marginal);

// multi-GPU (multigpu.cu)
void mg_destroy(vrgpu_ctx *ctx);
void mg_unique_id(char *out128);
void mg_init(vrgpu_ctx *ctx, int nranks, int rank, const char *id128);
int mg_nranks(const vrgpu_ctx *ctx);
int mg_rank(const vrgpu_ctx *ctx);
// What a multi-GPU search leaves for the host write-back of the coordinates the path rewrote (AdjustStructureForPeriod,
// search.cxx:393-396 / :856-895; the scaling round trip of the 6DFOF, :565-575): with `phase` the positions and velocities of
// every member of a 3D group (labels in pf3) changed in P, otherwise only the nmoved positions listed in midx / mxyz.
struct MgWriteback {
    DBuf<long long> pf3;
    DBuf<uint32_t> midx;
    DBuf<double> mxyz;
    size_t nmoved = 0;
    bool phase = false;
};
// wb == nullptr: P is left untouched (the resident variant restarts every run from it)
void mg_search(vrgpu_ctx *ctx, DeviceParticles &P, double slab_lo, double slab_hi, const vrgpu_search_opts *o, DBuf<long long> &pfof_out,
               long long &ngroups, MgWriteback *wb = nullptr);

} // namespace vrgpu
