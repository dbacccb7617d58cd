// velociraptor-stf_b200/csrc/ptree.cuh -- the kd-tree of ptree.cu as seen by its two users (potentials, local velocity density).
#pragma once
#include "groups.cuh"

#include <vector>

namespace vrgpu {

// Nodes are numbered level by level: the roots of the qualifying groups first (in group order), then the children of every
// splitting node of a level, left child at an even offset and right child = left + 1.  left == 0xffffffff marks a leaf.
struct PtTree {
    uint32_t nL = 0, nLG = 0, nnodes = 0;       // members in trees, number of trees (roots), nodes
    DBuf<uint32_t> slot_of;                     // compact member index -> slot in the group-major columns
    DBuf<uint32_t> perm;                        // tree position -> compact member index
    DBuf<uint32_t> nstart, nend, nleft;         // per node: tree positions [start, end), left child
    DBuf<uint32_t> rootidx;                     // tree position -> root node
    DBuf<double> rec;                           // the build's 32-byte records in tree order: x, y, z, (member, 0) -- 4 doubles per position
    DBuf<uint32_t> big_list, sub_list, sub_count; // nodes above 2048 particles (all levels); roots of the subtrees of at most 2048; sub_count[1] = nsub
    uint32_t nbig = 0, nsub = 0;
    std::vector<uint32_t> level_off;            // nodes of level l: [level_off[l], level_off[l + 1])
    std::vector<uint32_t> big_off;              // nodes above 2048 particles of level l: big_list[big_off[l], big_off[l + 1])
};

// NBody::KDTree::BuildNodes (NBodylib/src/KDTree/KDTree.cxx:438-507) with MedianPos's own arrangement (:304-337) for every
// group of gs with more than pp_threshold members.
bool pt_build(vrgpu_ctx *ctx, const GroupSet &gs, int bucket, long long pp_threshold, PtTree &t);

} // namespace vrgpu
