// tests/host/link2_emul.cpp -- serial emulation of the candidate-queue 3D linker (velociraptor-stf_b200/csrc/link2.cuh).
//
// Test infrastructure.  The per-thread functions of link2.cuh are compiled here by g++ (-ffp-contract=off) and run once
// per "thread" in a chosen order (forward, reverse or shuffled, to show that the result does not depend on the schedule);
// the surrounding steps (cell grid, sort, cell table, the exhaustive dense-pair pass) are plain host restatements of the
// device steps in fof3d.cu.  Output: the root cell of every particle, which tests/test_link2_host.py turns into
// canonical group ids and compares with the oracle's KDTree::FOF restatement.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <numeric>
#include <random>
#include <vector>

#include "link2.cuh"

using namespace vrgpu::link2;

namespace {

struct Order {
    std::vector<uint64_t> idx;
    Order(uint64_t n, int mode, uint32_t seed) : idx(n)
    {
        std::iota(idx.begin(), idx.end(), 0ull);
        if (mode == 1) std::reverse(idx.begin(), idx.end());
        else if (mode == 2) { std::mt19937_64 g(seed); std::shuffle(idx.begin(), idx.end(), g); }
    }
};

// the exhaustive test of k_link_queue / try_link_cells (fof3d.cu): every pair, both orientations across a periodic face
bool cells_link(const Params &P, uint32_t A, uint32_t B, uint32_t meta)
{
    const int sc = (int)(meta & 63u);
    const bool wrapped = sc != (1 | (1 << 2) | (1 << 4));
    const double shx = (double)((sc & 3) - 1) * P.period, shy = (double)(((sc >> 2) & 3) - 1) * P.period,
                 shz = (double)(((sc >> 4) & 3) - 1) * P.period;
    for (uint32_t i = P.cell_start[A]; i < P.cell_start[A + 1]; i++)
        for (uint32_t j = P.cell_start[B]; j < P.cell_start[B + 1]; j++) {
            const double ax = P.sx[i], ay = P.sy[i], az = P.sz[i], qx = P.sx[j], qy = P.sy[j], qz = P.sz[j];
            if (!wrapped) { if (pair_links(ax, ay, az, qx, qy, qz, P.b2)) return true; }
            else if (pair_links(ax + shx, ay + shy, az + shz, qx, qy, qz, P.b2) || pair_links(qx - shx, qy - shy, qz - shz, ax, ay, az, P.b2))
                return true;
        }
    return false;
}

} // namespace

// stats_out: [0] ncells, [1] candidates pass 1, [2] candidates pass 2, [3] dense pairs pass 1, [4] dense pairs pass 2,
//            [5] unions found by the inline trials, [6] unions found by the dense pass
extern "C" int link2_emul(const double *pos, long long n, double b, double period, int perx, int pery, int perz, int order_mode,
                          unsigned seed, int filter_pass2, long long *root_of_particle, long long *stats_out)
{
    for (int k = 0; k < 7; k++) stats_out[k] = 0;
    if (n == 0) return 0;
    // ---- grid (fof3d_link) ----
    const int periodic[3] = {period > 0 && perx, period > 0 && pery, period > 0 && perz};
    const double cmax = b / std::sqrt(3.0) * (1.0 - 1e-8);
    double lo[3], inv_c[3];
    int nc[3];
    for (int k = 0; k < 3; k++) {
        if (periodic[k]) {
            lo[k] = 0.0;
            nc[k] = (int)std::ceil(period / cmax);
            if (nc[k] < 8) return -1;
            inv_c[k] = (double)nc[k] / period;
        } else {
            double mn = pos[k], mx = pos[k];
            for (long long i = 1; i < n; i++) { mn = std::min(mn, pos[3 * i + k]); mx = std::max(mx, pos[3 * i + k]); }
            lo[k] = mn;
            nc[k] = (int)(std::floor((mx - mn) / cmax) + 1.0);
            inv_c[k] = 1.0 / cmax;
        }
    }
    // ---- keys, sort, cell table ----
    std::vector<uint64_t> key(n);
    for (long long i = 0; i < n; i++) {
        long long c[3];
        for (int k = 0; k < 3; k++) {
            c[k] = (long long)std::floor((pos[3 * i + k] - lo[k]) * inv_c[k]);
            c[k] = c[k] < 0 ? 0 : (c[k] >= nc[k] ? nc[k] - 1 : c[k]);
        }
        key[i] = ((uint64_t)c[2] * (uint64_t)nc[1] + (uint64_t)c[1]) * (uint64_t)nc[0] + (uint64_t)c[0];
    }
    std::vector<uint32_t> order(n);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t c) { return key[a] < key[c]; });
    std::vector<double> sx(n), sy(n), sz(n);
    std::vector<uint32_t> cell_of(n), cell_start;
    std::vector<uint64_t> cell_key;
    for (long long k = 0; k < n; k++) {
        const uint32_t i = order[k];
        sx[k] = pos[3 * i]; sy[k] = pos[3 * i + 1]; sz[k] = pos[3 * i + 2];
        if (k == 0 || key[i] != key[order[k - 1]]) { cell_start.push_back((uint32_t)k); cell_key.push_back(key[i]); }
        cell_of[k] = (uint32_t)cell_key.size() - 1;
    }
    const uint32_t ncells = (uint32_t)cell_key.size();
    cell_start.push_back((uint32_t)n);
    stats_out[0] = ncells;
    // ---- coordinates + row table ----
    const uint64_t nrows = (uint64_t)nc[1] * (uint64_t)nc[2];
    std::vector<RowRange> rows(nrows, RowRange{ROW_EMPTY, ROW_EMPTY});
    std::vector<uint32_t> cell_row(ncells), parent(ncells);
    std::vector<int> cell_ix(ncells);
    for (uint32_t c = 0; c < ncells; c++) coords_thread(cell_key.data(), ncells, (uint64_t)nc[0], c, cell_row.data(), cell_ix.data(), rows.data());
    std::iota(parent.begin(), parent.end(), 0u);

    Params P{};
    P.sx = sx.data(); P.sy = sy.data(); P.sz = sz.data();
    P.cell_start = cell_start.data(); P.cell_row = cell_row.data(); P.cell_ix = cell_ix.data(); P.rows = rows.data();
    P.parent = parent.data(); P.ncells = ncells;
    P.ncx = nc[0]; P.ncy = nc[1]; P.ncz = nc[2];
    P.period = period > 0 ? period : 0.0; P.b2 = b * b;
    P.perx = periodic[0]; P.pery = periodic[1]; P.perz = periodic[2];

    const uint32_t cap = (uint32_t)std::min<uint64_t>(16ull * ncells + 1024ull, 0x7fffffffull); // overflow is reported, not hidden
    std::vector<uint32_t> ca(cap), cb(cap), cm(cap), qa(cap), qb(cap), qm(cap);
    unsigned int ccount = 0, qcount = 0;
    Queue C{ca.data(), cb.data(), cm.data(), &ccount, cap};
    Queue Q{qa.data(), qb.data(), qm.data(), &qcount, cap};

    for (int reach = 1; reach <= 2; reach++) {
        ccount = 0; qcount = 0;
        const int nroles = roles_for(reach);
        {   // emit: grid (cells, roles)
            Order o((uint64_t)ncells * nroles, order_mode, seed + 11 * reach);
            if (reach == 2) for (uint64_t t : Order(ncells, order_mode, seed + 7).idx) compress_thread(P.parent, (uint32_t)t);
            for (uint64_t t : o.idx) emit_thread(P, (uint32_t)(t % ncells), (int)(t / ncells), reach, reach == 2 && filter_pass2, C);
        }
        if (ccount > C.capacity) return -2;
        stats_out[reach] = ccount;
        {   // pairs
            std::vector<uint32_t> before(parent);
            Order o(ccount, order_mode, seed + 101 * reach);
            for (uint64_t e : o.idx) pair_thread(P, (uint32_t)e, C, Q);
            for (uint32_t c = 0; c < ncells; c++) stats_out[5] += parent[c] != before[c] && before[c] == c;
        }
        if (qcount > Q.capacity) return -3;
        stats_out[2 + reach] = qcount;
        {   // dense pairs, exhaustive
            Order o(qcount, order_mode, seed + 1001 * reach);
            for (uint64_t e : o.idx) {
                const uint32_t A = Q.a[e], B = Q.b[e];
                if (find(P.parent, A) == find(P.parent, B)) continue;
                if (cells_link(P, A, B, Q.meta[e])) { unite(P.parent, A, B); stats_out[6]++; }
            }
        }
    }
    for (long long k = 0; k < n; k++) root_of_particle[order[k]] = (long long)find(P.parent, cell_of[k]);
    return 0;
}
