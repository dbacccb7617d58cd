#include "stages.cuh"
namespace vrgpu {
long long fof6d_device(vrgpu_ctx *, DeviceParticles &, const long long *, long long, double, double, int, long long, long long *, bool) { throw Error(VRGPU_EUNSUPPORTED, "fof6d not built yet"); }
void writeback_phase(vrgpu_ctx *, const vrgpu_view &, const DeviceParticles &, const long long *) { throw Error(VRGPU_EUNSUPPORTED, "nyi"); }
void potential_csr_host(vrgpu_ctx *, const DeviceParticles &, long long, const long long *, const long long *, double, double, double, int, long long, double *) { throw Error(VRGPU_EUNSUPPORTED, "nyi"); }
void unbind_device(vrgpu_ctx *, const DeviceParticles &, long long *, long long &, const vrgpu_unbind_opts &, int &, long long &) { throw Error(VRGPU_EUNSUPPORTED, "nyi"); }
}
