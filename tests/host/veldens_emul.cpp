// tests/host/veldens_emul.cpp -- serial emulation of the local velocity density (velociraptor-stf_b200/csrc/veldens.cuh).
//
// Test infrastructure.  The per-thread functions of veldens.cuh (leaf centres, neighbour search, velocity step + kernel
// sum) are compiled here by g++ (-ffp-contract=off) and run once per "thread"; the steps around them are plain host
// restatements of the device steps in veldens.cu: the tree (recursive here, with the reference's selection restated below:
// the node membership and order of the exact-ties build, which tests/test_gpu_tree.py pins on the device), tree-ordered columns,
// bounding boxes, the leaf list and the kernel table.  Queries run in batches like the device code.
// Output: densities in input order, which tests/test_veldens_host.py compares with the oracle's GetVelocityDensity.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#include <cstdint>

// The reference's median selection (NBody::KDTree::MedianPos, KDTree.cxx:304-337) as the literal loop over (value, index) pairs: the
// arrangement the library's kd-tree build reproduces on the device (csrc/ptree.cu: k_kd_big / k_kd_small), which
// tests/test_gpu_tree.py compares with the compiled reference through vrgpu_tree_order and tests/test_hoare_model.py pins on the CPU.
namespace emul_sel {
static inline uint32_t median_index(uint32_t start, uint32_t end) { return start + (end - start - 1u) / 2u; }
static inline void select_serial(double *val, uint32_t *idx, uint32_t k, uint32_t n)
{
    if (n == 0u) return;
    int left = 0, right = (int)n - 1;
    const int kk = (int)k;
    while (left < right) {
        const double x = val[kk];
        double wv = val[right]; val[right] = val[kk]; val[kk] = wv;
        uint32_t wi = idx[right]; idx[right] = idx[kk]; idx[kk] = wi;
        int i = left - 1, j = right;
        for (;;) {
            while (i < j) if (val[++i] >= x) break;
            while (i < j) if (val[--j] <= x) break;
            wv = val[i]; val[i] = val[j]; val[j] = wv;
            wi = idx[i]; idx[i] = idx[j]; idx[j] = wi;
            if (j <= i) break;
        }
        val[j] = val[i]; idx[j] = idx[i];
        val[i] = val[right]; idx[i] = idx[right];
        val[right] = wv; idx[right] = wi;
        if (i >= kk) right = i - 1;
        if (i <= kk) left = i + 1;
    }
}
} // namespace emul_sel
#include "veldens.cuh"

using namespace vrgpu::veldens;

namespace {

struct Tree {
    std::vector<uint32_t> nstart, nend, nleft, perm;
    const double *pos;
    int bucket;
};

// KDTree::BuildNodes (KDTree.cxx:438-507); children are allocated as a pair so that right = left + 1 as on the device
void build(Tree &t, uint32_t nd)
{
    const uint32_t s = t.nstart[nd], e = t.nend[nd];
    if (e - s <= (uint32_t)t.bucket) return;
    int d = 0;
    double best = 0;
    for (int k = 0; k < 3; k++) {
        double mn = t.pos[3 * (size_t)t.perm[s] + k], mx = mn;
        for (uint32_t p = s; p < e; p++) { mn = std::min(mn, t.pos[3 * (size_t)t.perm[p] + k]); mx = std::max(mx, t.pos[3 * (size_t)t.perm[p] + k]); }
        if (k == 0 || mx - mn > best) { best = mx - mn; d = k; }
    }
    std::vector<double> val(e - s);
    std::vector<uint32_t> idx(e - s);
    for (uint32_t i = 0; i < e - s; i++) { idx[i] = t.perm[s + i]; val[i] = t.pos[3 * (size_t)idx[i] + d]; }
    const uint32_t k = emul_sel::median_index(s, e);
    emul_sel::select_serial(val.data(), idx.data(), k - s, e - s);
    for (uint32_t i = 0; i < e - s; i++) t.perm[s + i] = idx[i];
    const uint32_t l = (uint32_t)t.nstart.size();
    t.nleft[nd] = l;
    t.nstart.push_back(s); t.nend.push_back(k + 1); t.nleft.push_back(NONE);
    t.nstart.push_back(k + 1); t.nend.push_back(e); t.nleft.push_back(NONE);
    build(t, l);
    build(t, l + 1);
}

void bbox(const Tree &t, uint32_t nd, const std::vector<double> &tx, const std::vector<double> &ty, const std::vector<double> &tz,
          std::vector<double> &bb)
{
    double *b = bb.data() + 6 * (size_t)nd;
    const uint32_t l = t.nleft[nd];
    if (l == NONE) {
        const uint32_t s = t.nstart[nd];
        b[0] = b[3] = tx[s]; b[1] = b[4] = ty[s]; b[2] = b[5] = tz[s];
        for (uint32_t p = s + 1; p < t.nend[nd]; p++) {
            b[0] = std::min(b[0], tx[p]); b[3] = std::max(b[3], tx[p]);
            b[1] = std::min(b[1], ty[p]); b[4] = std::max(b[4], ty[p]);
            b[2] = std::min(b[2], tz[p]); b[5] = std::max(b[5], tz[p]);
        }
        return;
    }
    bbox(t, l, tx, ty, tz, bb);
    bbox(t, l + 1, tx, ty, tz, bb);
    for (int k = 0; k < 3; k++) {
        b[k] = std::min(bb[6 * (size_t)l + k], bb[6 * (size_t)(l + 1) + k]);
        b[3 + k] = std::max(bb[6 * (size_t)l + 3 + k], bb[6 * (size_t)(l + 1) + 3 + k]);
    }
}

} // namespace

// stats_out: [0] leaves, [1] batches.  Returns 0, or < 0 for arguments the device call rejects as well.
// leaf_off (may be null; nleaves + 1 offsets): leaves of a tree the caller holds, the particles being in that tree's order.
extern "C" int veldens_emul(const double *pos, const double *vel, const int *type, long long n64, const long long *leaf_off, long long nleaves,
                            double period, int bsize, int nsearch, int nvel, int approx, unsigned batch, double *den_out, long long *stats_out)
{
    stats_out[0] = stats_out[1] = 0;
    if (n64 <= 0) return 0;
    if (nvel < 1 || nvel > MAXVEL || nsearch <= nvel || n64 - 1 < nsearch) return -1;
    const uint32_t n = (uint32_t)n64;
    Tree t;
    t.pos = pos;
    t.bucket = bsize;
    t.perm.resize(n);
    for (uint32_t i = 0; i < n; i++) t.perm[i] = i;
    t.nstart.push_back(0); t.nend.push_back(n); t.nleft.push_back(NONE);
    build(t, 0);
    const uint32_t nnodes = (uint32_t)t.nstart.size();

    std::vector<double> tx(n), ty(n), tz(n), tvx(n), tvy(n), tvz(n), den(n, 0.0), bb(6 * (size_t)nnodes);
    std::vector<int> ttype(n);
    for (uint32_t p = 0; p < n; p++) {
        const uint32_t i = t.perm[p];
        tx[p] = pos[3 * (size_t)i]; ty[p] = pos[3 * (size_t)i + 1]; tz[p] = pos[3 * (size_t)i + 2];
        tvx[p] = vel[3 * (size_t)i]; tvy[p] = vel[3 * (size_t)i + 1]; tvz[p] = vel[3 * (size_t)i + 2];
        ttype[p] = type ? type[i] : 1;
    }
    bbox(t, 0, tx, ty, tz, bb);
    std::vector<uint32_t> leaf;
    for (uint32_t nd = 0; nd < nnodes; nd++) if (t.nleft[nd] == NONE) leaf.push_back(nd);
    const bool caller_leaves = approx && leaf_off && nleaves > 0;
    const uint32_t nleaf = caller_leaves ? (uint32_t)nleaves : (uint32_t)leaf.size();
    stats_out[0] = nleaf;

    // KDTree::KernelConstruction (KDTree.cxx:592-631), ND = 3, Epanechnikov, 1000 entries -- as epanechnikov_table in veldens.cu
    const int kernres = 1000;
    std::vector<double> kernel(kernres);
    {
        const int ND = 3;
        double kernnorm = 1.0 / ND * std::pow(M_1_PI, ND / 2.) * std::tgamma(ND / 2. + 1.);
        kernnorm *= ND * (ND + 2.) * std::pow(0.5, ND + 1.);
        const double delta = 2.0 / (double)(kernres - 1);
        for (int i = 0; i < kernres; i++) {
            const double rh = i * delta / 1.0 * 0.5;
            kernel[i] = kernnorm * (rh <= 1.0 ? (1.0 - rh * rh) : 0.);
        }
    }

    std::vector<uint32_t> leaf_of(n), cnum(nleaf);
    std::vector<double> cm(3 * (size_t)nleaf);
    Params P{};
    P.tx = tx.data(); P.ty = ty.data(); P.tz = tz.data(); P.tvx = tvx.data(); P.tvy = tvy.data(); P.tvz = tvz.data();
    P.ttype = ttype.data(); P.n = n;
    P.nstart = t.nstart.data(); P.nend = t.nend.data(); P.nleft = t.nleft.data(); P.bb = bb.data(); P.root = 0;
    P.leaf = leaf.data(); P.leaf_of = leaf_of.data(); P.cm = cm.data(); P.cnum = cnum.data(); P.nleaf = nleaf;
    P.period = period; P.nsearch = (uint32_t)nsearch; P.nvel = (uint32_t)nvel; P.approx = approx ? 1 : 0;
    P.kernel = kernel.data(); P.kernres = kernres; P.den = den.data();
    if (caller_leaves) {
        std::vector<uint32_t> leaf_of_orig(n);
        CallerLeaves C{pos, type, leaf_off, nleaf, leaf_of_orig.data(), cm.data(), cnum.data()};
        for (uint32_t l = 0; l < nleaf; l++) caller_leaf_thread(C, l);
        for (uint32_t p = 0; p < n; p++) leaf_of[p] = leaf_of_orig[t.perm[p]];   // k_vd_map_leaf_of
    } else
        for (uint32_t l = 0; l < nleaf; l++) leaf_thread(P, l);

    const uint32_t nq = approx ? nleaf : n;
    const uint32_t nb = batch ? batch : nq;
    std::vector<double> heap_d2((size_t)nb * (nsearch + 1));
    std::vector<uint32_t> heap_id((size_t)nb * (nsearch + 1)), nn((size_t)nb * nsearch);
    P.heap_d2 = heap_d2.data(); P.heap_id = heap_id.data(); P.nn = nn.data();
    for (uint32_t q0 = 0; q0 < nq; q0 += nb) {
        P.q0 = q0;
        P.q1 = std::min(nq, q0 + nb);
        std::fill(nn.begin(), nn.end(), 0xdeadbeefu);            // stale results must never be read
        for (uint32_t q = P.q1; q-- > P.q0;) knn_thread(P, q);    // any order
        for (uint32_t j = 0; j < n; j++) density_thread(P, j);
        stats_out[1]++;
    }
    for (uint32_t p = 0; p < n; p++) den_out[t.perm[p]] = den[p];
    return 0;
}
