// velociraptor-stf_b200/csrc/veldens.cu -- local velocity density of one structure.
//
// Replaces GetVelocityDensity (src/localfield.cxx:18) as SearchSubSub calls it for the particles of one structure
// (search.cxx:2646, tree == NULL): the reference builds a physical kd-tree (bucket Options::Bsize, Epanechnikov kernel
// table of 1000 entries, period Options::p), runs one neighbour search per leaf (default) or per particle, and writes
// Particle::density.  Device stages:
//   tree      pt_build (ptree.cu: the reference's MedianPos arrangement): the leaves are the reference's
//   gather    tree-ordered position / velocity / type columns; bounding box of every node (leaves, then level by level up)
//   leaves    list of leaf nodes, centre and typed count of each (veldens::leaf_thread)
//   search    veldens::knn_thread per query, queries in batches bounded by the scratch of their queues
//   density   veldens::density_thread per particle of the batch
// The per-thread code lives in veldens.cuh and is shared with the host emulation (tests/host/veldens_emul.cpp).
#include "ptree.cuh"
#include "stages.cuh"
#include "veldens.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

namespace vrgpu {

namespace {

constexpr uint32_t NOCHILD = 0xffffffffu;

__global__ void k_vd_all_one(long long *__restrict__ a, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = 1;
}
__global__ void k_vd_gather(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
                            const double *__restrict__ vx, const double *__restrict__ vy, const double *__restrict__ vz,
                            const uint32_t *__restrict__ perm, const uint32_t *__restrict__ slot_of, const uint32_t *__restrict__ idx,
                            const int *__restrict__ type /* original order, may be null */, uint32_t n, double *__restrict__ tx,
                            double *__restrict__ ty, double *__restrict__ tz, double *__restrict__ tvx, double *__restrict__ tvy,
                            double *__restrict__ tvz, int *__restrict__ ttype)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t s = slot_of[perm[p]];
    tx[p] = x[s]; ty[p] = y[s]; tz[p] = z[s];
    tvx[p] = vx[s]; tvy[p] = vy[s]; tvz[p] = vz[s];
    ttype[p] = type ? type[idx[s]] : 1;
}
// bounding boxes: leaves from their particles, inner nodes from their children (one level per launch, bottom up)
__global__ void k_vd_bbox_leaf(const uint32_t *__restrict__ nstart, const uint32_t *__restrict__ nend, const uint32_t *__restrict__ nleft,
                               uint32_t nnodes, const double *__restrict__ tx, const double *__restrict__ ty, const double *__restrict__ tz,
                               double *__restrict__ bb)
{
    const uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= nnodes || nleft[nd] != NOCHILD) return;
    const uint32_t s = nstart[nd], e = nend[nd];
    double lo[3] = {tx[s], ty[s], tz[s]}, hi[3] = {tx[s], ty[s], tz[s]};
    for (uint32_t p = s + 1; p < e; p++) {
        lo[0] = fmin(lo[0], tx[p]); hi[0] = fmax(hi[0], tx[p]);
        lo[1] = fmin(lo[1], ty[p]); hi[1] = fmax(hi[1], ty[p]);
        lo[2] = fmin(lo[2], tz[p]); hi[2] = fmax(hi[2], tz[p]);
    }
    for (int k = 0; k < 3; k++) { bb[6 * (size_t)nd + k] = lo[k]; bb[6 * (size_t)nd + 3 + k] = hi[k]; }
}
__global__ void k_vd_bbox_inner(const uint32_t *__restrict__ nleft, uint32_t n0, uint32_t n1, double *__restrict__ bb)
{
    const uint32_t nd = n0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (nd >= n1) return;
    const uint32_t l = nleft[nd];
    if (l == NOCHILD) return;
    for (int k = 0; k < 3; k++) {
        bb[6 * (size_t)nd + k] = fmin(bb[6 * (size_t)l + k], bb[6 * (size_t)(l + 1) + k]);
        bb[6 * (size_t)nd + 3 + k] = fmax(bb[6 * (size_t)l + 3 + k], bb[6 * (size_t)(l + 1) + 3 + k]);
    }
}
__global__ void k_vd_leaf_flag(const uint32_t *__restrict__ nleft, uint32_t nnodes, uint32_t *__restrict__ flags)
{
    const uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd < nnodes) flags[nd] = nleft[nd] == NOCHILD ? 1u : 0u;
}
__global__ void k_vd_leaf_list(const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl, uint32_t nnodes, uint32_t *__restrict__ leaf)
{
    const uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
    if (nd < nnodes && flags[nd]) leaf[excl[nd]] = nd;
}
__global__ void k_vd_leaves(veldens::Params P)
{
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < P.nleaf) veldens::leaf_thread(P, l);
}
__global__ void k_vd_caller_leaves(veldens::CallerLeaves C)
{
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < C.nleaf) veldens::caller_leaf_thread(C, l);
}
// tree position -> caller's leaf
__global__ void k_vd_map_leaf_of(const uint32_t *__restrict__ leaf_of_orig, const uint32_t *__restrict__ perm, const uint32_t *__restrict__ slot_of,
                                 const uint32_t *__restrict__ idx, uint32_t n, uint32_t *__restrict__ leaf_of)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) leaf_of[p] = leaf_of_orig[idx[slot_of[perm[p]]]];
}
__global__ void __launch_bounds__(128) k_vd_knn(veldens::Params P)
{
    const uint32_t q = P.q0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (q < P.q1) veldens::knn_thread(P, q);
}
__global__ void __launch_bounds__(128) k_vd_density(veldens::Params P)
{
    // per-leaf search: any particle can belong to a leaf of the batch; per-particle search: the batch IS a range of particles
    const uint32_t j = (P.approx ? 0u : P.q0) + blockIdx.x * blockDim.x + threadIdx.x;
    if (j < (P.approx ? P.n : P.q1)) veldens::density_thread(P, j);
}
__global__ void k_vd_scatter(const double *__restrict__ den, const uint32_t *__restrict__ perm, const uint32_t *__restrict__ slot_of,
                             const uint32_t *__restrict__ idx, uint32_t n, double *__restrict__ out)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) out[idx[slot_of[perm[p]]]] = den[p];
}

} // namespace

// KDTree::KernelConstruction (KDTree.cxx:592-631) for ND = 3 and the Epanechnikov kernel, evaluated on the host with the
// C library the reference would use on this machine
static std::vector<double> epanechnikov_table(int kernres)
{
    const int ND = 3;
    double kernnorm = 1.0 / ND * std::pow(M_1_PI, ND / 2.) * std::tgamma(ND / 2. + 1.);
    kernnorm *= ND * (ND + 2.) * std::pow(0.5, ND + 1.);
    const double delta = 2.0 / (double)(kernres - 1);
    std::vector<double> k((size_t)kernres);
    for (int i = 0; i < kernres; i++) {
        const double rh = i * delta / 1.0 * 0.5;
        k[(size_t)i] = kernnorm * (rh <= 1.0 ? (1.0 - rh * rh) : 0.);
    }
    return k;
}

void velocity_density_device(vrgpu_ctx *ctx, const DeviceParticles &P, const int *type_dev, const long long *leaf_off_dev, long long ncaller_leaves,
                             double period, int bsize, int nsearch, int nvel, int approx, double *den_dev)
{
    const int T = 256;
    VR_REQUIRE(P.has_vel, VRGPU_EINVAL, "the velocity density needs velocities");
    VR_REQUIRE(bsize >= 1, VRGPU_EINVAL, "Bsize must be >= 1");
    VR_REQUIRE(nvel >= 1 && nvel <= veldens::MAXVEL, VRGPU_EUNSUPPORTED, "Nvel must be in [1, 64]");
    VR_REQUIRE(nsearch > nvel, VRGPU_EINVAL, "Nsearch must exceed Nvel");
    VR_REQUIRE(period >= 0 && std::isfinite(period), VRGPU_EINVAL, "period must be >= 0");
    if (P.n == 0) return;
    VR_REQUIRE((long long)P.n - 1 >= nsearch, VRGPU_EINVAL, "fewer particles than Nsearch + 1");
    VR_REQUIRE(P.n < 0x7fffffffull, VRGPU_EUNSUPPORTED, "more than 2^31 particles in one structure");
    const uint32_t n = (uint32_t)P.n;
    VR_CUDA(cudaMemsetAsync(den_dev, 0, sizeof(double) * P.n, ctx->stream));

    // ---- one group holding every particle, its tree with the reference's leaves ----
    GroupSet gs;
    {
        DBuf<long long> one(ctx, P.n);
        VR_LAUNCH(ctx, k_vd_all_one, grid_for(P.n, T), T, 0, one.p, P.n);
        build_groupset(ctx, one.p, P.n, 1, gs);
    }
    gather_groupset(ctx, P, gs, true, false);
    PtTree t;
    {
        KScope k(ctx, "veldens.tree");
        const bool ok = pt_build(ctx, gs, bsize, 0, t);
        t.rec.reset();
        VR_REQUIRE(ok && t.nL == n && t.nLG == 1, VRGPU_ECUDA, "velocity density: tree build failed");
    }
    const uint32_t nnodes = t.nnodes;

    // ---- tree-ordered columns, bounding boxes, leaves ----
    DBuf<double> tx(ctx, n), ty(ctx, n), tz(ctx, n), tvx(ctx, n), tvy(ctx, n), tvz(ctx, n), bb(ctx, 6 * (size_t)nnodes), den(ctx, n);
    DBuf<int> ttype(ctx, n);
    DBuf<uint32_t> leaf, leaf_of(ctx, n);
    uint32_t nleaf = 0;
    {
        KScope k(ctx, "veldens.prepare");
        VR_LAUNCH(ctx, k_vd_gather, grid_for(n, T), T, 0, gs.x.p, gs.y.p, gs.z.p, gs.vx.p, gs.vy.p, gs.vz.p, t.perm.p, t.slot_of.p, gs.idx.p,
                  type_dev, n, tx.p, ty.p, tz.p, tvx.p, tvy.p, tvz.p, ttype.p);
        VR_LAUNCH(ctx, k_vd_bbox_leaf, grid_for(nnodes, T), T, 0, t.nstart.p, t.nend.p, t.nleft.p, nnodes, tx.p, ty.p, tz.p, bb.p);
        for (size_t l = t.level_off.size() - 1; l-- > 0;) {
            const uint32_t n0 = t.level_off[l], n1 = t.level_off[l + 1];
            if (n1 > n0) VR_LAUNCH(ctx, k_vd_bbox_inner, grid_for(n1 - n0, T), T, 0, t.nleft.p, n0, n1, bb.p);
        }
        DBuf<uint32_t> flags(ctx, nnodes), excl(ctx, nnodes), total(ctx, 1);
        VR_LAUNCH(ctx, k_vd_leaf_flag, grid_for(nnodes, T), T, 0, t.nleft.p, nnodes, flags.p);
        exclusive_scan_u32(ctx, flags.p, excl.p, nnodes, total.p);
        VR_CUDA(cudaMemcpyAsync(&nleaf, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        leaf.alloc(ctx, nleaf);
        VR_LAUNCH(ctx, k_vd_leaf_list, grid_for(nnodes, T), T, 0, flags.p, excl.p, nnodes, leaf.p);
    }
    // leaves of the caller's tree instead of our own (per-leaf search only; the search itself still walks our tree)
    const bool caller_leaves = approx && leaf_off_dev && ncaller_leaves > 0;
    if (caller_leaves) nleaf = (uint32_t)ncaller_leaves;
    DBuf<double> cm(ctx, 3 * (size_t)nleaf), kernel(ctx, 1000);
    DBuf<uint32_t> cnum(ctx, nleaf);
    const std::vector<double> ktab = epanechnikov_table(1000);
    VR_CUDA(cudaMemcpyAsync(kernel.p, ktab.data(), sizeof(double) * ktab.size(), cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream)); // ktab is a local

    veldens::Params V{};
    V.tx = tx.p; V.ty = ty.p; V.tz = tz.p; V.tvx = tvx.p; V.tvy = tvy.p; V.tvz = tvz.p; V.ttype = ttype.p; V.n = n;
    V.nstart = t.nstart.p; V.nend = t.nend.p; V.nleft = t.nleft.p; V.bb = bb.p; V.root = 0;
    V.leaf = leaf.p; V.leaf_of = leaf_of.p; V.cm = cm.p; V.cnum = cnum.p; V.nleaf = nleaf;
    V.period = period; V.nsearch = (uint32_t)nsearch; V.nvel = (uint32_t)nvel; V.approx = approx ? 1 : 0;
    V.kernel = kernel.p; V.kernres = 1000; V.den = den.p;
    if (caller_leaves) {
        DBuf<uint32_t> leaf_of_orig(ctx, n);
        veldens::CallerLeaves C{P.pos.p, type_dev, leaf_off_dev, nleaf, leaf_of_orig.p, cm.p, cnum.p};
        VR_LAUNCH(ctx, k_vd_caller_leaves, grid_for(nleaf, T), T, 0, C);
        VR_LAUNCH(ctx, k_vd_map_leaf_of, grid_for(n, T), T, 0, leaf_of_orig.p, t.perm.p, t.slot_of.p, gs.idx.p, n, leaf_of.p);
    } else VR_LAUNCH(ctx, k_vd_leaves, grid_for(nleaf, T), T, 0, V);

    // ---- queries in batches: the queues take (Nsearch + 1) * 12 bytes and the result Nsearch * 4 bytes per query ----
    const uint32_t nq = approx ? nleaf : n;
    const size_t per_query = (size_t)(nsearch + 1) * 12 + (size_t)nsearch * 4;
    const uint32_t batch = (uint32_t)std::max<size_t>(1024, std::min<size_t>(nq, ((size_t)1 << 30) / per_query));
    DBuf<double> heap_d2(ctx, (size_t)batch * (nsearch + 1));
    DBuf<uint32_t> heap_id(ctx, (size_t)batch * (nsearch + 1)), nn(ctx, (size_t)batch * nsearch);
    V.heap_d2 = heap_d2.p; V.heap_id = heap_id.p; V.nn = nn.p;
    DBuf<unsigned int> ovf(ctx, 1);
    VR_CUDA(cudaMemsetAsync(ovf.p, 0, sizeof(unsigned int), ctx->stream));
    V.overflow = ovf.p;
    for (uint32_t q0 = 0; q0 < nq; q0 += batch) {
        V.q0 = q0;
        V.q1 = (uint32_t)std::min<size_t>(nq, (size_t)q0 + batch);
        { KScope k(ctx, "veldens.search"); VR_LAUNCH(ctx, k_vd_knn, grid_for(V.q1 - V.q0, 128), 128, 0, V); }
        { KScope k(ctx, "veldens.density"); VR_LAUNCH(ctx, k_vd_density, grid_for(approx ? n : V.q1 - V.q0, 128), 128, 0, V); }
    }
    VR_LAUNCH(ctx, k_vd_scatter, grid_for(n, T), T, 0, den.p, t.perm.p, t.slot_of.p, gs.idx.p, n, den_dev);
    unsigned int hovf = 0;
    VR_CUDA(cudaMemcpyAsync(&hovf, ovf.p, sizeof(hovf), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    VR_REQUIRE(hovf == 0, VRGPU_EUNSUPPORTED, "velocity density: the neighbour search ran out of stack (tree deeper than 47 levels)");
}

} // namespace vrgpu
