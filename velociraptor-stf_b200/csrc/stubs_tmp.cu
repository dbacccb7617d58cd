#include "stages.cuh"
#include "groups.cuh"
namespace vrgpu {
void mg_destroy(vrgpu_ctx *) {}
void potential_tree(vrgpu_ctx *, const GroupSet &, double, double, double, int, long long, double *) { throw Error(VRGPU_EUNSUPPORTED, "tree potential not built yet"); }
}
