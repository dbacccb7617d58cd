// tests/host/link3_emul.cpp -- host emulation of the 3D cell linker's neighbour enumeration (velociraptor-stf_b200/csrc/link3.cuh).
//
// Test infrastructure.  link3.cuh is plain integer code shared with the kernels of fof3d.cu; here g++ compiles it and
//   link3_check_tables  compares the (local index, block) candidate masks with a brute-force walk over the 62 half-space
//                       offsets of a cell,
//   link3_fof           runs the whole linker serially (grid as fof3d_link builds it, block-major sort, block tables,
//                       lookup_blocks / cell_of_bit per cell, the reference's pair predicate, union-find) and returns one
//                       label per particle -- tests/test_link3_host.py compares the components with the oracle's 3DFOF.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

#include "link3.cuh"

using namespace vrgpu::link3;

extern "C" long long link3_check_tables()
{
    std::vector<uint64_t> tab(TABLE_WORDS);
    build_table(tab.data());
    long long bad = 0;
    for (unsigned loc = 0; loc < 64; loc++) {
        const int lx = loc & 3, ly = (loc >> 2) & 3, lz = loc >> 4;
        uint64_t want[2][8];
        std::memset(want, 0, sizeof(want));
        for (int dz = -2; dz <= 2; dz++)
            for (int dy = -2; dy <= 2; dy++)
                for (int dx = -2; dx <= 2; dx++) {
                    const bool half = dz > 0 || (dz == 0 && dy > 0) || (dz == 0 && dy == 0 && dx > 0);
                    if (!half) continue;
                    const int m = std::max(std::abs(dx), std::max(std::abs(dy), std::abs(dz)));
                    const int x = lx + dx, y = ly + dy, z = lz + dz; // relative to the own block's origin
                    const int ox = x < 0 ? -1 : (x > 3 ? 1 : 0), oy = y < 0 ? -1 : (y > 3 ? 1 : 0), oz = z < 0 ? -1 : (z > 3 ? 1 : 0);
                    // the other block along an axis is the lower one for local 0, 1 and the upper one for 2, 3
                    if ((ox == -1 && lx >= 2) || (ox == 1 && lx < 2) || (oy == -1 && ly >= 2) || (oy == 1 && ly < 2) || (oz == -1 && lz >= 2) ||
                        (oz == 1 && lz < 2)) { bad += 1000000; continue; }
                    const unsigned j = (ox ? 1u : 0u) | (oy ? 2u : 0u) | (oz ? 4u : 0u);
                    const unsigned bit = (unsigned)((x - 4 * ox) + 4 * (y - 4 * oy) + 16 * (z - 4 * oz));
                    want[m <= 1 ? 0 : 1][j] |= 1ull << bit;
                    int ddx, ddy, ddz;
                    cell_offset(loc, j, bit, ddx, ddy, ddz);
                    if (ddx != dx || ddy != dy || ddz != dz) bad++;
                }
        for (int pass = 0; pass < 2; pass++)
            for (unsigned j = 0; j < 8; j++)
                if (tab[(pass * 64 + loc) * 8 + j] != want[pass][j]) bad++;
    }
    return bad;
}

namespace {
struct UF {
    std::vector<uint32_t> p;
    uint32_t find(uint32_t x) { while (p[x] != x) { p[x] = p[p[x]]; x = p[x]; } return x; }
    void unite(uint32_t a, uint32_t b) { a = find(a); b = find(b); if (a == b) return; if (a < b) std::swap(a, b); p[a] = b; }
};
inline bool pair_links(double ax, double ay, double az, double bx, double by, double bz, double b2)
{
    const double dx = ax - bx, dy = ay - by, dz = az - bz;
    const double d2 = (dx * dx + dy * dy) + dz * dz; // compiled with -ffp-contract=off
    return d2 < b2;
}
} // namespace

// pos: n x 3; periodic[3]: which axes wrap at `period` (period <= 0: none).  labels_out[i] = smallest particle index of i's
// component.  order_mode 0 / 1: cells visited forward / backward.  Returns the number of non-empty cells, < 0 on error.
extern "C" long long link3_fof(const double *pos, long long n, double b, double period, const int *periodic, int order_mode, long long *labels_out,
                               long long *stats_out /* [0] cells, [1] blocks, [2] candidates, [3] pair tests */)
{
    if (n <= 0) return 0;
    const double cmax = b / std::sqrt(3.0) * (1.0 - 1e-8);
    Geom g{};
    double lo[3], invc[3];
    int nc[3], per[3];
    for (int k = 0; k < 3; k++) {
        per[k] = period > 0 && periodic[k];
        if (per[k]) {
            lo[k] = 0.0;
            long long m = (long long)std::ceil(period / cmax);
            if (m < 8) return -2;
            m = (m + 3) / 4 * 4; // whole blocks, so that the grid wraps block on block
            // cells three apart must not link: 2 c >= b.  Small boxes (below ~15 b) may have no multiple of 4 in the valid
            // range of cell counts; the library takes its hash-table path for those (fof3d.cu)
            if (2.0 * (period / (double)m) < b * (1.0 + 1e-6)) return -5;
            nc[k] = (int)m;
            invc[k] = (double)nc[k] / period;
        } else {
            double mn = 1e300, mx = -1e300;
            for (long long i = 0; i < n; i++) { mn = std::min(mn, pos[3 * i + k]); mx = std::max(mx, pos[3 * i + k]); }
            lo[k] = mn;
            nc[k] = (int)(std::floor((mx - mn) / cmax) + 1.0);
            invc[k] = 1.0 / cmax;
        }
    }
    g.ncx = nc[0]; g.ncy = nc[1]; g.ncz = nc[2];
    g.nbx = (nc[0] + 3) / 4; g.nby = (nc[1] + 3) / 4; g.nbz = (nc[2] + 3) / 4;
    g.perx = per[0]; g.pery = per[1]; g.perz = per[2];
    std::vector<uint64_t> key(n);
    for (long long i = 0; i < n; i++) {
        int c[3];
        for (int k = 0; k < 3; k++) {
            long long v = (long long)std::floor((pos[3 * i + k] - lo[k]) * invc[k]);
            v = v < 0 ? 0 : (v >= nc[k] ? nc[k] - 1 : v);
            c[k] = (int)v;
        }
        key[i] = cell_key(g, c[0], c[1], c[2]);
        int x, y, z;
        key_cell(g, key[i], x, y, z);
        if (x != c[0] || y != c[1] || z != c[2]) return -3;
    }
    std::vector<uint32_t> order(n);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t c) { return key[a] < key[c]; });
    // cells, blocks
    std::vector<uint32_t> cell_start, cell_blk, blk_cell0;
    std::vector<uint8_t> cell_loc;
    std::vector<unsigned long long> blk_occ;
    std::vector<uint64_t> blk_xyz, blk_lin;
    for (long long k = 0; k < n; k++) {
        const uint64_t ky = key[order[k]];
        if (k == 0 || ky != key[order[k - 1]]) {
            if (k == 0 || (ky >> 6) != (key[order[k - 1]] >> 6)) {
                int x, y, z;
                key_cell(g, ky, x, y, z);
                blk_xyz.push_back(pack_xyz(x >> 2, y >> 2, z >> 2));
                blk_lin.push_back(ky >> 6);
                blk_cell0.push_back((uint32_t)cell_start.size());
                blk_occ.push_back(0ull);
            }
            blk_occ.back() |= 1ull << (ky & 63u);
            cell_blk.push_back((uint32_t)blk_occ.size() - 1u);
            cell_loc.push_back((uint8_t)(ky & 63u));
            cell_start.push_back((uint32_t)k);
        }
    }
    const uint32_t ncells = (uint32_t)cell_start.size();
    cell_start.push_back((uint32_t)n);
    const uint64_t nblk_total = (uint64_t)g.nbx * g.nby * g.nbz;
    const size_t nwords = (size_t)(nblk_total / 64 + 1);
    if (nwords > ((size_t)1 << 28)) return -4;
    std::vector<unsigned long long> bitmap(nwords, 0ull);
    for (uint64_t l : blk_lin) bitmap[l >> 6] |= 1ull << (l & 63u);
    std::vector<uint32_t> prefix(nwords);
    uint32_t run = 0;
    for (size_t w = 0; w < nwords; w++) { prefix[w] = run; run += (uint32_t)__builtin_popcountll(bitmap[w]); }
    Tables T{bitmap.data(), prefix.data(), blk_occ.data(), blk_cell0.data()};
    std::vector<uint64_t> tab(TABLE_WORDS);
    build_table(tab.data());

    UF uf;
    uf.p.resize(ncells);
    std::iota(uf.p.begin(), uf.p.end(), 0u);
    const double b2 = b * b;
    long long ncand = 0, ntests = 0;
    for (int pass = 0; pass < 2; pass++)
        for (uint32_t cc = 0; cc < ncells; cc++) {
            const uint32_t c = order_mode ? ncells - 1 - cc : cc;
            int bx, by, bz;
            unpack_xyz(blk_xyz[cell_blk[c]], bx, by, bz);
            const unsigned loc = cell_loc[c];
            uint32_t rank[8];
            unsigned long long occ[8], codes;
            lookup_blocks(g, T, bx, by, bz, loc, rank, occ, codes);
            if (rank[0] != cell_blk[c]) return -6;
            { int x2, y2, z2; unsigned l2; unpack_cell(pack_cell(bx, by, bz, loc), x2, y2, z2, l2); if (x2 != bx || y2 != by || z2 != bz || l2 != loc) return -7; }
            for (unsigned j = 0; j < 8; j++) {
                unsigned long long cand = occ[j] & tab[(pass * 64 + loc) * 8 + j];
                while (cand) {
                    const unsigned bit = (unsigned)__builtin_ctzll(cand);
                    cand &= cand - 1;
                    const uint32_t B = cell_of_bit(T, rank[j], occ[j], bit);
                    ncand++;
                    if (uf.find(c) == uf.find(B)) continue;
                    const int sc = (int)((codes >> (6 * j)) & 63u);
                    const double shx = (double)((sc & 3) - 1) * period, shy = (double)(((sc >> 2) & 3) - 1) * period,
                                 shz = (double)(((sc >> 4) & 3) - 1) * period;
                    const bool wrapped = sc != (1 | (1 << 2) | (1 << 4));
                    bool linked = false;
                    for (uint32_t ia = cell_start[c]; ia < cell_start[c + 1] && !linked; ia++)
                        for (uint32_t ib = cell_start[B]; ib < cell_start[B + 1] && !linked; ib++) {
                            const double *a = pos + 3 * (size_t)order[ia], *q = pos + 3 * (size_t)order[ib];
                            ntests++;
                            if (!wrapped) linked = pair_links(a[0], a[1], a[2], q[0], q[1], q[2], b2);
                            else
                                linked = pair_links(a[0] + shx, a[1] + shy, a[2] + shz, q[0], q[1], q[2], b2) ||
                                         pair_links(q[0] - shx, q[1] - shy, q[2] - shz, a[0], a[1], a[2], b2);
                        }
                    if (linked) uf.unite(c, B);
                }
            }
        }
    // label = smallest particle index of the component
    std::vector<long long> minidx(ncells, -1);
    std::vector<uint32_t> cell_of(n);
    for (uint32_t c = 0; c < ncells; c++)
        for (uint32_t k = cell_start[c]; k < cell_start[c + 1]; k++) {
            cell_of[k] = c;
            const uint32_t r = uf.find(c);
            if (minidx[r] < 0 || (long long)order[k] < minidx[r]) minidx[r] = order[k];
        }
    for (long long k = 0; k < n; k++) labels_out[order[k]] = minidx[uf.find(cell_of[k])];
    if (stats_out) { stats_out[0] = ncells; stats_out[1] = (long long)blk_occ.size(); stats_out[2] = ncand; stats_out[3] = ntests; }
    return ncells;
}
