// velociraptor-stf_b200/csrc/link2.cuh -- candidate-queue variant of the 3D cell linker (VRGPU_LINK_VARIANT=1, fof3d.cu).
//
// Same job as k_link_thread (fof3d.cu): unite the non-empty fine cells of the FOF grid whose particles satisfy the
// reference's pair predicate (KDLeafNode.cxx:556-558, DistFunc.h:27-32, KDFOF.cxx:14; periodic images as
// KDSplitNode.cxx:1500-1539).  The stall samples of k_link_thread (profiles/r01h_stall_analysis.md) show 3-6 of 32 lanes
// active in its inner loops and 52 % of the samples behind dependent loads: a thread that owns a cell with neighbours
// does the root look-ups, the particle loads and the pair trials while the lanes of the (mostly isolated) neighbouring
// cells idle.  Here the work is split so that every phase is one short, uniform thread:
//   emit  : one thread per (cell, neighbour ROW).  A warp has 32 consecutive cells and ONE row offset, so the row-table
//           loads of a warp fall into one or two sectors and most threads leave after that single load.  A thread whose
//           row holds cells inside the x-interval appends (cell, neighbour cell, image code) to a candidate queue.
//   pairs : one thread per candidate: two root look-ups, at most 16 pair trials, union; cell pairs with more than 16
//           particle pairs that the trial left undecided go to the dense-pair queue finished by k_link_queue (a warp per
//           pair, exhaustive).
// Pass 1 covers the 13 adjacent half-space cells (reach 1), pass 2 the 49 second-shell ones; between them every cell is
// pointed at its root and pass 2 drops candidates that are already in one set when they are generated (at 128^3: 2.9 M
// candidates in pass 1, 8.9 M possible and 0.09 M emitted in pass 2).  VRGPU_LINK_VARIANT=2 emits pass 2 unfiltered.
//
// Every function here is per-thread code without warp intrinsics (the only exception, the warp-aggregated queue
// append, has a trivial serial stand-in), so the SAME source is compiled by g++ into the serial emulation that
// tests/test_link2_host.py checks against the oracle: tests/host/link2_emul.cpp.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define L2_HD __device__ __forceinline__
#define L2_HOSTDEV __host__ __device__ __forceinline__
#else
#define L2_HD static inline
#define L2_HOSTDEV static inline
#endif

namespace vrgpu {
namespace link2 {

constexpr uint32_t ROW_EMPTY = 0xffffffffu;

struct alignas(8) RowRange {
    uint32_t first, last; // cells [first, last) of one (iy, iz) row, ordered by ix; first == ROW_EMPTY: no cell
};

struct Params {
    const double *sx, *sy, *sz;   // positions in cell order
    const uint32_t *cell_start;   // ncells + 1
    const uint32_t *cell_row;     // ncells: iz * ncy + iy
    const int *cell_ix;           // ncells
    const RowRange *rows;         // ncy * ncz
    uint32_t *parent;             // union-find over cells
    uint32_t ncells;
    int ncx, ncy, ncz;
    double period, b2;
    int perx, pery, perz;
};

// meta: image code sc (6 bits: (s+1) per axis, 2 bits each) | (ddx+2)<<6 | (ddy+2)<<9 | (ddz+2)<<12 -- the layout
// k_link_queue reads
struct Queue {
    uint32_t *a, *b, *meta;
    unsigned int *count;
    uint32_t capacity;
};

#if defined(__CUDACC__)
L2_HD uint32_t ld_parent(const uint32_t *p, uint32_t i) { return ((const volatile uint32_t *)p)[i]; }
L2_HD uint32_t cas_u32(uint32_t *a, uint32_t cmp, uint32_t v) { return atomicCAS(a, cmp, v); }
L2_HD uint32_t ld_u32(const uint32_t *p) { return __ldg(p); }
L2_HD int ld_i32(const int *p) { return __ldg(p); }
L2_HD RowRange ld_row(const RowRange *p)
{
    const uint2 v = __ldg(reinterpret_cast<const uint2 *>(p));
    return RowRange{v.x, v.y};
}
L2_HD double fsub(double a, double b) { return __dsub_rn(a, b); }
L2_HD double fadd(double a, double b) { return __dadd_rn(a, b); }
L2_HD double fmul(double a, double b) { return __dmul_rn(a, b); }
// one atomicAdd per converged group of lanes
L2_HD unsigned int queue_slot(unsigned int *count)
{
    const unsigned m = __activemask();
    unsigned lane;
    asm("mov.u32 %0, %%laneid;" : "=r"(lane));
    const int leader = __ffs(m) - 1;
    unsigned int base = 0;
    if ((int)lane == leader) base = atomicAdd(count, (unsigned int)__popc(m));
    base = __shfl_sync(m, base, leader);
    return base + (unsigned int)__popc(m & ((1u << lane) - 1u));
}
#else
// serial emulation (compiled with -ffp-contract=off: plain IEEE operations, no fused multiply-add)
L2_HD uint32_t ld_parent(const uint32_t *p, uint32_t i) { return p[i]; }
L2_HD uint32_t cas_u32(uint32_t *a, uint32_t cmp, uint32_t v)
{
    const uint32_t old = *a;
    if (old == cmp) *a = v;
    return old;
}
L2_HD uint32_t ld_u32(const uint32_t *p) { return *p; }
L2_HD int ld_i32(const int *p) { return *p; }
L2_HD RowRange ld_row(const RowRange *p) { return *p; }
L2_HD double fsub(double a, double b) { return a - b; }
L2_HD double fadd(double a, double b) { return a + b; }
L2_HD double fmul(double a, double b) { return a * b; }
L2_HD unsigned int queue_slot(unsigned int *count) { return (*count)++; }
#endif

L2_HD bool pair_links(double ax, double ay, double az, double bx, double by, double bz, double b2)
{
    const double dx = fsub(ax, bx), dy = fsub(ay, by), dz = fsub(az, bz);
    const double d2 = fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz));
    return d2 < b2;
}

// roots carry the smallest index of their set (same rules as uf_find / uf_union in ufhash.cuh)
L2_HD uint32_t find(uint32_t *parent, uint32_t x)
{
    uint32_t p = ld_parent(parent, x);
    while (p != x) {
        const uint32_t gp = ld_parent(parent, p);
        if (gp != p) parent[x] = gp; // path halving; a stale value is still an ancestor
        x = p;
        p = gp;
    }
    return x;
}

L2_HD void unite(uint32_t *parent, uint32_t a, uint32_t b)
{
    for (;;) {
        a = find(parent, a);
        b = find(parent, b);
        if (a == b) return;
        if (a < b) { const uint32_t t = a; a = b; b = t; }
        if (cas_u32(&parent[a], a, b) == a) return;
    }
}

L2_HOSTDEV int roles_for(int reach) { return reach == 1 ? 5 : 13; }

// (dy, dz) of a role.  reach 1: own row, (1,0), (-1..1, 1).  reach 2: own row, (1..2, 0), (-2..2, 1), (-2..2, 2).
L2_HD void role_offsets(int reach, int role, int &dy, int &dz)
{
    if (role == 0) { dy = 0; dz = 0; }
    else if (reach == 1) {
        if (role == 1) { dy = 1; dz = 0; }
        else { dy = role - 3; dz = 1; }
    } else {
        if (role < 3) { dy = role; dz = 0; }
        else if (role < 8) { dy = role - 5; dz = 1; }
        else { dy = role - 10; dz = 2; }
    }
}

// fills cell_row / cell_ix and the row table (rows[] preset to 0xff bytes)
L2_HD void coords_thread(const uint64_t *cell_key, uint32_t ncells, uint64_t ncx, uint32_t c, uint32_t *cell_row, int *cell_ix, RowRange *rows)
{
    const uint64_t key = cell_key[c];
    const uint64_t r = key / ncx;
    cell_row[c] = (uint32_t)r;
    cell_ix[c] = (int)(key - r * ncx);
    if (c == 0 || cell_key[c - 1] / ncx != r) rows[r].first = c;
    if (c == ncells - 1 || cell_key[c + 1] / ncx != r) rows[r].last = c + 1;
}

// between the passes: point every cell at its root, so that the filter of pass 2 needs two loads per look-up
L2_HD void compress_thread(uint32_t *parent, uint32_t c)
{
    const uint32_t r = find(parent, c);
    if (r != c) parent[c] = r;
}

// one (cell, role): append the neighbour cells of that row to the candidate queue; `filter` drops neighbours that are
// already in the set of `cell`
L2_HD void emit_thread(const Params &P, const uint32_t cell, const int role, const int reach, const bool filter, const Queue C)
{
    int dy, dz;
    role_offsets(reach, role, dy, dz);
    const uint32_t row = ld_u32(&P.cell_row[cell]);
    const int ix = ld_i32(&P.cell_ix[cell]);
    const int iz = (int)(row / (uint32_t)P.ncy);
    const int iy = (int)(row - (uint32_t)iz * (uint32_t)P.ncy);
    int ny = iy + dy, nz = iz + dz, sy = 0, sz = 0;
    if (ny < 0) { if (!P.pery) return; ny += P.ncy; sy = 1; }
    else if (ny >= P.ncy) { if (!P.pery) return; ny -= P.ncy; sy = -1; }
    if (nz >= P.ncz) { if (!P.perz) return; nz -= P.ncz; sz = -1; }
    const uint32_t r = (uint32_t)nz * (uint32_t)P.ncy + (uint32_t)ny;
    const RowRange rr = ld_row(&P.rows[r]);
    const uint32_t rf = rr.first, rl = rr.last;
    if (rf == ROW_EMPTY) return;
    const int x0 = role == 0 ? ix + 1 : ix - reach, x1 = ix + reach; // unwrapped x-interval
    // cells at Chebyshev distance <= 1 were handled by the reach-1 pass
    const bool near_row = reach == 2 && dy >= -1 && dy <= 1 && dz <= 1;
    uint32_t ra = ROW_EMPTY; // root of `cell`, looked up when the first candidate appears (only when filtering)
    for (int seg = 0; seg < 2; seg++) {
        int xlo = x0, xhi = x1, xoff = 0, sx = 0;
        if (seg == 0) { xlo = xlo > 0 ? xlo : 0; xhi = xhi < P.ncx - 1 ? xhi : P.ncx - 1; }
        else {
            if (!P.perx) break;
            if (x0 < 0) { xhi = (x1 < -1 ? x1 : -1) + P.ncx; xlo = x0 + P.ncx; xoff = -P.ncx; sx = 1; }
            else if (x1 >= P.ncx) { xlo = (x0 > P.ncx ? x0 : P.ncx) - P.ncx; xhi = x1 - P.ncx; xoff = P.ncx; sx = -1; }
            else break;
        }
        if (xlo > xhi) continue;
        uint32_t c;
        if (role == 0 && seg == 0) c = cell + 1;
        else {
            uint32_t lo = rf, hi = rl;
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if (ld_i32(&P.cell_ix[mid]) < xlo) lo = mid + 1; else hi = mid;
            }
            c = lo;
        }
        const uint32_t sc = (uint32_t)((sx + 1) | ((sy + 1) << 2) | ((sz + 1) << 4));
        for (; c < rl; c++) {
            const int bx = ld_i32(&P.cell_ix[c]);
            if (bx > xhi) break;
            const int ddx = bx + xoff - ix;
            if (near_row && ddx >= -1 && ddx <= 1) continue;
            if (filter) {
                if (ra == ROW_EMPTY) ra = find(P.parent, cell);
                if (find(P.parent, c) == ra) continue;
            }
            const unsigned int slot = queue_slot(C.count);
            if (slot < C.capacity) {
                C.a[slot] = cell;
                C.b[slot] = c;
                C.meta[slot] = sc | ((uint32_t)(ddx + 2) << 6) | ((uint32_t)(dy + 2) << 9) | ((uint32_t)(dz + 2) << 12);
            }
        }
    }
}

// one candidate: root check, inline pair trial (all pairs when there are at most 16, else the first 4 x 4), union or
// hand-over to the dense-pair queue
L2_HD void pair_thread(const Params &P, const uint32_t e, const Queue C, const Queue Q)
{
    const uint32_t cell = C.a[e], B = C.b[e], m = C.meta[e];
    if (find(P.parent, cell) == find(P.parent, B)) return;
    const uint32_t a0 = P.cell_start[cell], na = P.cell_start[cell + 1] - a0;
    const uint32_t b0 = P.cell_start[B], nb = P.cell_start[B + 1] - b0;
    const bool few = (uint64_t)na * (uint64_t)nb <= 16u;
    const uint32_t ta = few ? na : (na < 4u ? na : 4u), tb = few ? nb : (nb < 4u ? nb : 4u);
    const int sc = (int)(m & 63u);
    const bool wrapped = sc != (1 | (1 << 2) | (1 << 4));
    const double shx = (double)((sc & 3) - 1) * P.period, shy = (double)(((sc >> 2) & 3) - 1) * P.period,
                 shz = (double)(((sc >> 4) & 3) - 1) * P.period;
    bool linked = false;
    for (uint32_t i = 0; i < ta && !linked; i++) {
        const double ax = P.sx[a0 + i], ay = P.sy[a0 + i], az = P.sz[a0 + i];
        for (uint32_t j = 0; j < tb && !linked; j++) {
            const double qx = P.sx[b0 + j], qy = P.sy[b0 + j], qz = P.sz[b0 + j];
            if (!wrapped) linked = pair_links(ax, ay, az, qx, qy, qz, P.b2);
            else
                linked = pair_links(fadd(ax, shx), fadd(ay, shy), fadd(az, shz), qx, qy, qz, P.b2) ||
                         pair_links(fsub(qx, shx), fsub(qy, shy), fsub(qz, shz), ax, ay, az, P.b2);
        }
    }
    if (linked) unite(P.parent, cell, B);
    else if (!few) {
        const unsigned int slot = queue_slot(Q.count);
        if (slot < Q.capacity) {
            Q.a[slot] = cell;
            Q.b[slot] = B;
            Q.meta[slot] = m;
        }
    }
}

} // namespace link2
} // namespace vrgpu
