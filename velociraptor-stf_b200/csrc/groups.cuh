// velociraptor-stf_b200/csrc/groups.cuh -- group-major ("regrouped") device layout shared by 6DFOF / potentials / unbinding.
//
// Replaces the reference's AoS qsort by group id (src/search.cxx:422-440) and the per-group deep copies of
// BuildPartList / BuildPGList (src/buildandsortarrays.cxx:26-43,134): grouped particles are compacted, stably
// radix-sorted by group id (members keep ascending particle index, like pglist) and their columns gathered
// into SoA arrays indexed by SLOT.
#pragma once
#include "fof.cuh"

namespace vrgpu {

struct GroupSet {
    size_t ntot = 0;        // number of grouped particles
    long long ngroups = 0;
    DBuf<uint32_t> idx;     // slot -> original particle index
    DBuf<uint32_t> gid;     // slot -> group id (1..G)
    DBuf<uint32_t> off;     // G+2 entries: group g occupies slots [off[g], off[g+1]); off[0] = off[1] = 0
    // gathered columns (filled on request)
    DBuf<double> x, y, z, vx, vy, vz, m;
};

// pfof_dev: device array (original order) with ids in [0, ngroups]
void build_groupset(vrgpu_ctx *ctx, const long long *pfof_dev, size_t n, long long ngroups, GroupSet &gs);
void gather_groupset(vrgpu_ctx *ctx, const DeviceParticles &P, GroupSet &gs, bool want_vel, bool want_mass);

// Order candidate groups by (size desc, min pid asc) and hand out ids 1..G.  `root[c] == c` marks candidate c,
// gsize / gminpid are indexed by candidate; gid_of[c] receives the new id (0 for non-candidates / too small).
uint32_t order_groups(vrgpu_ctx *ctx, const uint32_t *root, const uint32_t *gsize, const unsigned long long *gminpid,
                      uint32_t ncand, uint32_t minsize, uint32_t *gid_of);

// direct-sum potentials for every slot of gs whose group size is <= pp_threshold (or all when pp_threshold < 0);
// phi[slot] written for those slots only.  Returns the number of ordered pair interactions evaluated.
unsigned long long potential_direct(vrgpu_ctx *ctx, const GroupSet &gs, double G, double eps, long long pp_threshold, double *phi);
// kd-tree monopole potentials (Potential(), src/unbind.cxx:1607) for every group with size > pp_threshold >= 0
void potential_tree(vrgpu_ctx *ctx, const GroupSet &gs, double G, double eps, double theta, int bucket, long long pp_threshold,
                    double *phi);

// the particle order NBody::KDTree::BuildNodes leaves in every group (bucket = leaf size): out[slot] = offset inside its group of
// the member that sits at that tree position
void tree_order_device(vrgpu_ctx *ctx, const GroupSet &gs, int bucket, long long *out_dev);

// BuildNumInGroup + BuildPGList (src/buildandsortarrays.cxx:11-16, 26-43) from device labels: numingroup_out[0..ngroups],
// index_out = members of group 1, 2, ... in ascending particle index; returns the number of grouped particles
long long pglist_device(vrgpu_ctx *ctx, const long long *pfof_dev, size_t n, long long ngroups, long long *numingroup_out,
                        long long *index_out);

} // namespace vrgpu
