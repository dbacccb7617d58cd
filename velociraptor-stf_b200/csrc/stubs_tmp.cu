#include "stages.cuh"
#include "groups.cuh"
namespace vrgpu {
long long fof6d_device(vrgpu_ctx *, DeviceParticles &, const long long *, long long, double, double, int, long long, long long *, bool) { throw Error(VRGPU_EUNSUPPORTED, "fof6d not built yet"); }
void writeback_phase(vrgpu_ctx *, const vrgpu_view &, const DeviceParticles &, const long long *) { throw Error(VRGPU_EUNSUPPORTED, "nyi"); }
void mg_destroy(vrgpu_ctx *) {}
void potential_tree(vrgpu_ctx *, const GroupSet &, double, double, double, int, long long, double *) { throw Error(VRGPU_EUNSUPPORTED, "tree potential not built yet"); }
}
