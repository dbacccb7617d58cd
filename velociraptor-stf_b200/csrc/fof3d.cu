// velociraptor-stf_b200/csrc/fof3d.cu -- 3D friends-of-friends on the device.
//
// Replaces NBody::KDTree::FOF (NBodylib/src/KDTree/KDFOF.cxx:12-129) + the leaf/split-node ball searches
// (KDLeafNode.cxx:518-573, KDSplitNode.cxx:865-891, 1496-1540) as called by SearchFullSet (src/search.cxx:203).
//
// FOF groups are the transitive closure of the pair predicate, so the kd-tree BFS is not reproduced; instead
//   1. particles are binned into FINE cells of edge c < b/sqrt(3): two particles in one cell are always linked,
//      so the union-find runs over non-empty CELLS, not particles;
//   2. particles are radix-sorted by cell key (sort.cu), positions gathered into cell order (SoA, coalesced);
//   3. cells are grouped into 4 x 4 x 4 blocks (the sort key is block-major): a bitmap over the blocks with a per-word rank and a
//      64-bit occupancy mask per non-empty block turn "which of my 62 half-space neighbour cells exist" into two independent
//      look-ups and a table AND per neighbouring block (link3.cuh).  One thread per cell enumerates its candidates; the
//      candidates of a warp go through a per-warp queue in shared memory, so the root checks and pair trials always run on
//      full warps; cell pairs the short trial leaves undecided go to a warp-per-pair kernel (exhaustive test, fp64 predicate
//      of the reference); lock-free atomicCAS union with path halving over CELLS;
//   4. roots are flattened, group sizes / min pid accumulated per root, groups >= minsize ordered by
//      (size desc, min pid asc) and pfof written in the caller's particle order.
// Pair predicate: d2 = dx*dx + dy*dy + dz*dz summed left to right in fp64 WITHOUT fma, link iff d2 < b*b
// (KDLeafNode.cxx:556-558, DistFunc.h:27-32, KDFOF.cxx:14).  Periodic images: the target is shifted by +-period
// before the subtraction (KDSplitNode.cxx:1500-1539); either particle of a pair can be the BFS target in the
// reference, so both orientations are evaluated.
#include "fof.cuh"
#include "groups.cuh"
#include "ufhash.cuh"
#include "link3.cuh"

#include <algorithm>
#include <cmath>

namespace vrgpu {

struct LinkParams {
    const double *sx, *sy, *sz;
    const uint32_t *cell_start;
    const uint64_t *cell_key;
    uint32_t ncells;
    uint32_t *parent;
    const uint64_t *hkeys;
    const uint32_t *hvals;
    uint32_t hmask;
    int ncx, ncy, ncz;
    double cx, cy, cz;
    double lox, loy, loz;
    double period;
    double b2;
    int perx, pery, perz; // per-axis periodicity (a slab of a multi-GPU run is open along x)
    link3::Geom g;        // blocks of 4 x 4 x 4 cells; cell_key = (block, local), link3.cuh
};

__constant__ signed char c_offsets[62][3];

// ---------------------------------------------------------------------------------------------------
// binning
// ---------------------------------------------------------------------------------------------------
__global__ void k_cell_keys(const double *__restrict__ pos, size_t n, Grid g, link3::Geom geom, uint64_t *__restrict__ keys,
                            uint32_t *__restrict__ vals, unsigned int *__restrict__ nbad)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double x = pos[3 * i], y = pos[3 * i + 1], z = pos[3 * i + 2];
    bool bad = !(x == x) || !(y == y) || !(z == z);
    if (g.periodic[0]) bad = bad || x < 0 || x > g.period;
    if (g.periodic[1]) bad = bad || y < 0 || y > g.period;
    if (g.periodic[2]) bad = bad || z < 0 || z > g.period;
    if (bad) atomicAdd(nbad, 1u);
    long long ix = (long long)floor((x - g.lo[0]) * g.inv_c[0]);
    long long iy = (long long)floor((y - g.lo[1]) * g.inv_c[1]);
    long long iz = (long long)floor((z - g.lo[2]) * g.inv_c[2]);
    ix = ix < 0 ? 0 : (ix >= g.nc[0] ? g.nc[0] - 1 : ix);
    iy = iy < 0 ? 0 : (iy >= g.nc[1] ? g.nc[1] - 1 : iy);
    iz = iz < 0 ? 0 : (iz >= g.nc[2] ? g.nc[2] - 1 : iz);
    keys[i] = link3::cell_key(geom, (int)ix, (int)iy, (int)iz);
    vals[i] = (uint32_t)i;
}

__global__ void k_minmax(const double *__restrict__ pos, size_t n, double *__restrict__ out /*lo[3],hi[3]*/)
{
    // small helper for open boxes: per-block reduction + atomics on ordered-int representation is overkill here;
    // positions are finite doubles, use atomicMin/Max on the sortable 64-bit image.
    __shared__ double smin[3][32], smax[3][32];
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        for (int k = 0; k < 3; k++) {
            double v = pos[3 * i + k];
            mn[k] = fmin(mn[k], v);
            mx[k] = fmax(mx[k], v);
        }
    unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o; o >>= 1) {
            mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
        if (lane == 0) { smin[k][warp] = mn[k]; smax[k][warp] = mx[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        int k = threadIdx.x;
        double a = 1e300, b = -1e300;
        for (unsigned w = 0; w < (blockDim.x >> 5); w++) { a = fmin(a, smin[k][w]); b = fmax(b, smax[k][w]); }
        // ordered-int atomics: for doubles, map to sortable unsigned
        auto enc = [](double d) { unsigned long long u = (unsigned long long)__double_as_longlong(d); return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull); };
        atomicMin((unsigned long long *)&out[k], enc(a));
        atomicMax((unsigned long long *)&out[3 + k], enc(b));
    }
}

__global__ void k_gather_pos(const double *__restrict__ pos, const uint32_t *__restrict__ order, size_t n, double *__restrict__ sx,
                             double *__restrict__ sy, double *__restrict__ sz)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    size_t i = order[k];
    sx[k] = pos[3 * i];
    sy[k] = pos[3 * i + 1];
    sz[k] = pos[3 * i + 2];
}

// head flags of cells (low word) and of blocks (high word); one 64-bit scan numbers both
__global__ void k_flag_heads(const uint64_t *__restrict__ keys, size_t n, unsigned long long *__restrict__ flags)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint64_t key = keys[k];
    const uint64_t prev = k ? keys[k - 1] : ~key;
    flags[k] = (key != prev ? 1ull : 0ull) | ((key >> 6) != (prev >> 6) ? (1ull << 32) : 0ull);
}

struct BlockArrays {
    unsigned long long *cell_rec;  // cell -> block coordinates + local index (link3::pack_cell)
    uint32_t *blk_cell0;           // block rank -> first cell
    unsigned long long *blk_occ;   // block rank -> occupancy of its 64 cells (zeroed before)
    unsigned long long *bitmap;    // all blocks of the grid, one bit each (zeroed before; nullptr: hash path)
};

__global__ void k_fill_cells(const uint64_t *__restrict__ keys, const unsigned long long *__restrict__ excl, size_t n, link3::Geom geom,
                             uint32_t *__restrict__ cell_of, uint32_t *__restrict__ cell_start, uint64_t *__restrict__ cell_key, uint32_t ncells,
                             BlockArrays B)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint64_t key = keys[k];
    const uint64_t prev = k ? keys[k - 1] : ~key;
    const bool chead = key != prev, bhead = (key >> 6) != (prev >> 6);
    const unsigned long long e = excl[k];
    const uint32_t c = (uint32_t)(e & 0xffffffffull) - (chead ? 0u : 1u);
    cell_of[k] = c;
    if (chead) {
        cell_start[c] = (uint32_t)k;
        cell_key[c] = key;
        if (B.bitmap) {
            const uint32_t r = (uint32_t)(e >> 32) - (bhead ? 0u : 1u);
            int ix, iy, iz;
            link3::key_cell(geom, key, ix, iy, iz);
            B.cell_rec[c] = link3::pack_cell(ix >> 2, iy >> 2, iz >> 2, (unsigned)(key & 63u));
            atomicOr(&B.blk_occ[r], 1ull << (key & 63u));
            if (bhead) {
                B.blk_cell0[r] = c;
                const uint64_t lin = key >> 6;
                atomicOr(&B.bitmap[lin >> 6], 1ull << (lin & 63u));
            }
        }
    }
    if (k == n - 1) cell_start[ncells] = (uint32_t)n;
}

__global__ void k_bitmap_counts(const unsigned long long *__restrict__ bitmap, size_t nwords, uint32_t *__restrict__ cnt)
{
    size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w < nwords) cnt[w] = (uint32_t)__popcll(bitmap[w]);
}

__device__ __forceinline__ bool pair_links(double ax, double ay, double az, double bx, double by, double bz, double b2)
{
    double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by), dz = __dsub_rn(az, bz);
    double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
    return d2 < b2;
}

// Try to unite cell `cell` (particles [a0,a1), grid coordinates ix,iy,iz) with neighbour cell B reached through the cell
// offset (ddx,ddy,ddz); sc packs the periodic image code (s+1 per axis, 2 bits each).  Whole warp cooperates; stops at the
// first linking pair.
__device__ __forceinline__ void try_link_cells(const LinkParams &P, const uint32_t cell, const uint32_t a0, const uint32_t a1, const int ix,
                                               const int iy, const int iz, const uint32_t B, const int ddx, const int ddy, const int ddz,
                                               const int sc, const unsigned lane)
{
    const uint32_t na = a1 - a0;
    uint32_t ra = 0, rb = 0;
    if (lane == 0) { ra = uf_find(P.parent, cell); rb = uf_find(P.parent, B); }
    ra = __shfl_sync(0xffffffffu, ra, 0);
    rb = __shfl_sync(0xffffffffu, rb, 0);
    if (ra == rb) return;
    const double shx = (double)((sc & 3) - 1) * P.period, shy = (double)(((sc >> 2) & 3) - 1) * P.period,
                 shz = (double)(((sc >> 4) & 3) - 1) * P.period;
    const bool wrapped = sc != (1 | (1 << 2) | (1 << 4));
    const uint32_t b0 = P.cell_start[B], b1 = P.cell_start[B + 1];
    const uint32_t nbp = b1 - b0;
    bool linked = false;
    const uint64_t npairs = (uint64_t)na * nbp;
    if (npairs <= 4096) {
        for (uint32_t p0 = 0; p0 < (uint32_t)npairs && !linked; p0 += 32) {
            uint32_t p = p0 + lane;
            bool hit = false;
            if (p < (uint32_t)npairs) {
                uint32_t ia = a0 + p / nbp, ib = b0 + p % nbp;
                double ax = P.sx[ia], ay = P.sy[ia], az = P.sz[ia];
                double bx = P.sx[ib], by = P.sy[ib], bz = P.sz[ib];
                if (!wrapped) hit = pair_links(ax, ay, az, bx, by, bz, P.b2);
                else
                    hit = pair_links(__dadd_rn(ax, shx), __dadd_rn(ay, shy), __dadd_rn(az, shz), bx, by, bz, P.b2) ||
                          pair_links(__dsub_rn(bx, shx), __dsub_rn(by, shy), __dsub_rn(bz, shz), ax, ay, az, P.b2);
            }
            linked = __any_sync(0xffffffffu, hit);
        }
    } else {
        // big cell pair: only particles of A whose distance to B's box is below b can link
        const double blox = P.lox + (double)(ix + ddx) * P.cx, bloy = P.loy + (double)(iy + ddy) * P.cy,
                     bloz = P.loz + (double)(iz + ddz) * P.cz;
        const double margin = 1e-6 * P.cx;
        const double b2m = P.b2 * (1.0 + 1e-6);
        for (uint32_t abase = a0; abase < a1 && !linked; abase += 32) {
            uint32_t ia = abase + lane;
            double ax = 0, ay = 0, az = 0;
            bool near = false;
            if (ia < a1) {
                ax = P.sx[ia]; ay = P.sy[ia]; az = P.sz[ia];
                double gx = ddx > 0 ? blox - ax : (ddx < 0 ? ax - (blox + P.cx) : 0.0);
                double gy = ddy > 0 ? bloy - ay : (ddy < 0 ? ay - (bloy + P.cy) : 0.0);
                double gz = ddz > 0 ? bloz - az : (ddz < 0 ? az - (bloz + P.cz) : 0.0);
                gx = fmax(gx - margin, 0.0); gy = fmax(gy - margin, 0.0); gz = fmax(gz - margin, 0.0);
                near = gx * gx + gy * gy + gz * gz < b2m;
            }
            unsigned nm = __ballot_sync(0xffffffffu, near);
            while (nm && !linked) {
                const int s = __ffs(nm) - 1;
                nm &= nm - 1;
                const double tx = __shfl_sync(0xffffffffu, ax, s), ty = __shfl_sync(0xffffffffu, ay, s),
                             tz = __shfl_sync(0xffffffffu, az, s);
                const double txs = __dadd_rn(tx, shx), tys = __dadd_rn(ty, shy), tzs = __dadd_rn(tz, shz);
                for (uint32_t bb = b0; bb < b1 && !linked; bb += 32) {
                    uint32_t ib = bb + lane;
                    bool hit = false;
                    if (ib < b1) {
                        double bx = P.sx[ib], by = P.sy[ib], bz = P.sz[ib];
                        if (!wrapped) hit = pair_links(tx, ty, tz, bx, by, bz, P.b2);
                        else
                            hit = pair_links(txs, tys, tzs, bx, by, bz, P.b2) ||
                                  pair_links(__dsub_rn(bx, shx), __dsub_rn(by, shy), __dsub_rn(bz, shz), tx, ty, tz, P.b2);
                    }
                    linked = __any_sync(0xffffffffu, hit);
                }
            }
        }
    }
    if (linked && lane == 0) uf_union(P.parent, cell, B);
    __syncwarp();
}

// Hash-table path (grids too large for the block bitmap, periodic boxes too small for whole blocks): one warp per cell probes
// its half-space neighbours in a cell-key hash table.  off_begin/off_end select the slice of the 62 half-space offsets handled by this launch:
// the 13 face/edge/corner neighbours first, the 49 second-shell ones afterwards (by then most are already united).
__global__ void __launch_bounds__(256) k_link_cells(LinkParams P, int off_begin, int off_end)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t warps_total = (gridDim.x * blockDim.x) >> 5;
    const uint32_t warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (uint32_t cell = warp_global; cell < P.ncells; cell += warps_total) {
        int ix, iy, iz;
        link3::key_cell(P.g, P.cell_key[cell], ix, iy, iz);
        const uint32_t a0 = P.cell_start[cell], a1 = P.cell_start[cell + 1];
        const uint32_t na = a1 - a0;
        for (int obase = off_begin; obase < off_end; obase += 32) {
            const int o = obase + (int)lane;
            int nb = -1;
            int scode = 0; // (sx+1) | (sy+1)<<2 | (sz+1)<<4 | dxyz packed above
            int dx = 0, dy = 0, dz = 0;
            if (o < off_end) {
                dx = c_offsets[o][0]; dy = c_offsets[o][1]; dz = c_offsets[o][2];
                int nx = ix + dx, ny = iy + dy, nz = iz + dz;
                int sx = 0, sy = 0, sz = 0;
                bool ok = true;
                if (nx < 0) { if (P.perx) { nx += P.ncx; sx = 1; } else ok = false; }
                else if (nx >= P.ncx) { if (P.perx) { nx -= P.ncx; sx = -1; } else ok = false; }
                if (ny < 0) { if (P.pery) { ny += P.ncy; sy = 1; } else ok = false; }
                else if (ny >= P.ncy) { if (P.pery) { ny -= P.ncy; sy = -1; } else ok = false; }
                if (nz < 0) { if (P.perz) { nz += P.ncz; sz = 1; } else ok = false; }
                else if (nz >= P.ncz) { if (P.perz) { nz -= P.ncz; sz = -1; } else ok = false; }
                if (ok) {
                    const uint64_t nkey = link3::cell_key(P.g, nx, ny, nz);
                    nb = hash_lookup(P.hkeys, P.hvals, P.hmask, nkey);
                    scode = (sx + 1) | ((sy + 1) << 2) | ((sz + 1) << 4);
                }
            }
            unsigned found = __ballot_sync(0xffffffffu, nb >= 0);
            while (found) {
                const int src = __ffs(found) - 1;
                found &= found - 1;
                const uint32_t B = (uint32_t)__shfl_sync(0xffffffffu, nb, src);
                const int sc = __shfl_sync(0xffffffffu, scode, src);
                const int ddx = __shfl_sync(0xffffffffu, dx, src), ddy = __shfl_sync(0xffffffffu, dy, src),
                          ddz = __shfl_sync(0xffffffffu, dz, src);
                try_link_cells(P, cell, a0, a1, ix, iy, iz, B, ddx, ddy, ddz, sc, lane);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// block linker (link3.cuh).  Measured and replaced (profiles/r01h_stall_analysis.md, r02a_probe.md): a thread per cell that
// walks a dense (iy, iz) row table with binary searches -- 13 to 25 dependent look-up chains per cell, 9 of 32 lanes active.
// ---------------------------------------------------------------------------------------------------
struct LinkQueue {
    uint32_t *a, *b, *meta;     // meta: sc (6 bits) | (ddx+2)<<6 | (ddy+2)<<9 | (ddz+2)<<12
    unsigned int *count;
    uint32_t capacity;
};

struct BlockParams {
    link3::Tables T;
    const unsigned long long *cell_rec;
    const unsigned long long *table; // link3::build_table, TABLE_WORDS entries
};

// one candidate: are the cells A and B in one set already?  If not, a short trial of particle pairs (all of them when there are
// at most 16, else the first 4 x 4); a link unites the sets, an undecided dense pair goes to the queue of the warp-per-pair
// kernel (or is finished here when the queue is full).  cmeta: image code (6 bits) | block j << 6 | bit << 9 | local index << 15.
__device__ __forceinline__ void link_candidate(const LinkParams &P, const uint32_t A, const uint32_t B, const uint32_t cmeta, const LinkQueue &Q)
{
    // both root chases side by side, the cell ranges fetched while they run
    volatile uint32_t *vp = P.parent;
    uint32_t ra = A, rb = B, pa = vp[A], pb = vp[B];
    const uint32_t a0 = __ldg(P.cell_start + A), a1 = __ldg(P.cell_start + A + 1);
    const uint32_t b0 = __ldg(P.cell_start + B), b1 = __ldg(P.cell_start + B + 1);
    while (pa != ra || pb != rb) {
        const uint32_t ga = pa != ra ? vp[pa] : pa, gb = pb != rb ? vp[pb] : pb;
        if (pa != ra) { if (ga != pa) P.parent[ra] = ga; ra = pa; pa = ga; } // path halving; a stale value is still an ancestor
        if (pb != rb) { if (gb != pb) P.parent[rb] = gb; rb = pb; pb = gb; }
    }
    if (ra == rb) return;
    const uint32_t na = a1 - a0, nb = b1 - b0;
    const int sc = (int)(cmeta & 63u);
    const bool wrapped = sc != (1 | (1 << 2) | (1 << 4));
    const double shx = (double)((sc & 3) - 1) * P.period, shy = (double)(((sc >> 2) & 3) - 1) * P.period,
                 shz = (double)(((sc >> 4) & 3) - 1) * P.period;
    const bool few = na * nb <= 16u;
    uint32_t ta = few ? na : min(na, 4u), tb = few ? nb : min(nb, 4u);
    bool linked = false;
    for (int attempt = 0; attempt < 2 && !linked; attempt++) {
        for (uint32_t i = 0; i < ta && !linked; i++) {
            const double ax = P.sx[a0 + i], ay = P.sy[a0 + i], az = P.sz[a0 + i];
            for (uint32_t j = 0; j < tb && !linked; j++) {
                const double qx = P.sx[b0 + j], qy = P.sy[b0 + j], qz = P.sz[b0 + j];
                if (!wrapped) linked = pair_links(ax, ay, az, qx, qy, qz, P.b2);
                else
                    linked = pair_links(__dadd_rn(ax, shx), __dadd_rn(ay, shy), __dadd_rn(az, shz), qx, qy, qz, P.b2) ||
                             pair_links(__dsub_rn(qx, shx), __dsub_rn(qy, shy), __dsub_rn(qz, shz), ax, ay, az, P.b2);
            }
        }
        if (linked || few) break;
        const unsigned int slot = atomicAdd(Q.count, 1u);
        if (slot < Q.capacity) {
            int ddx, ddy, ddz;
            link3::cell_offset((cmeta >> 15) & 63u, (cmeta >> 6) & 7u, (cmeta >> 9) & 63u, ddx, ddy, ddz);
            Q.a[slot] = A;
            Q.b[slot] = B;
            Q.meta[slot] = (uint32_t)sc | ((uint32_t)(ddx + 2) << 6) | ((uint32_t)(ddy + 2) << 9) | ((uint32_t)(ddz + 2) << 12);
            break;
        }
        ta = na; tb = nb; // queue full (pathological density): this thread tests every pair itself
    }
    if (linked) uf_union(P.parent, A, B);
}

// after pass 0: the common root of a block's cells (0xffffffff when they differ).  Pass 1 skips a neighbouring block whose cells all
// sit in the cell's own set with ONE look-up instead of one per candidate cell -- inside a halo that is nearly every block.
__global__ void k_block_roots(const uint32_t *__restrict__ snap, const uint32_t *__restrict__ blk_cell0, uint32_t nblocks, uint32_t ncells,
                              uint32_t *__restrict__ blk_root)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nblocks) return;
    const uint32_t c0 = blk_cell0[r], c1 = r + 1 < nblocks ? blk_cell0[r + 1] : ncells;
    const uint32_t root = snap[c0];
    bool same = true;
    for (uint32_t c = c0 + 1; c < c1; c++) same = same && snap[c] == root;
    blk_root[r] = same ? root : 0xffffffffu;
}

// PASS 0: the 13 adjacent half-space cells of every cell (this unites most of a halo).  PASS 1: the 49 cells of the second shell,
// after every cell's root has been written to `snap`: a candidate whose snapshot root equals the cell's is dropped at once
// (two independent loads instead of two pointer chases) -- inside a halo that is nearly all of them.
constexpr int LB_THREADS = 256, LB_WARPS = LB_THREADS / 32, LB_POP = 4, LB_QCAP = 32 * (LB_POP + 1);
template <int PASS>
__global__ void __launch_bounds__(LB_THREADS, 4) k_link_blocks(LinkParams P, BlockParams Bp, const uint32_t *__restrict__ snap,
                                                               const uint32_t *__restrict__ blk_root, uint8_t *__restrict__ need2, LinkQueue Q)
{
    __shared__ unsigned long long s_tab[(PASS == 0 ? 2 : 1) * 64 * 8]; // pass 0 also looks at the second shell: who needs pass 1?
    __shared__ uint32_t s_qa[LB_WARPS][LB_QCAP], s_qb[LB_WARPS][LB_QCAP], s_qm[LB_WARPS][LB_QCAP];
    const unsigned tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int i = tid; i < (PASS == 0 ? 2 : 1) * 64 * 8; i += LB_THREADS) s_tab[i] = Bp.table[PASS * 64 * 8 + i];
    __syncthreads();
    const uint32_t cell = blockIdx.x * LB_THREADS + tid;
    bool valid = cell < P.ncells;
    if (PASS == 1) {
        valid = valid && need2[cell] != 0;
        if (!__any_sync(0xffffffffu, valid)) return; // no cell of this warp has a neighbour in the second shell
    }
    unsigned loc = 0;
    unsigned long long codes = 0;
    uint32_t rank[8], myroot = 0;
    unsigned long long occ[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { rank[j] = 0; occ[j] = 0; }
    if (valid) {
        int bx, by, bz;
        link3::unpack_cell(__ldg(Bp.cell_rec + cell), bx, by, bz, loc);
        link3::lookup_blocks(P.g, Bp.T, bx, by, bz, loc, rank, occ, codes);
        if (PASS == 1) myroot = __ldg(snap + cell);
    }
    // which of the 8 blocks hold candidates for some lane of the warp?
    unsigned jmask = 0;
    bool later = false;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        unsigned long long c = occ[j] & s_tab[loc * 8 + j];
        if (PASS == 1 && c != 0 && __ldg(blk_root + rank[j]) == myroot) { c = 0; occ[j] = 0; } // the whole block is in this cell's set already
        if (PASS == 0) later = later || (occ[j] & s_tab[64 * 8 + loc * 8 + j]) != 0;
        if (__any_sync(0xffffffffu, c != 0)) jmask |= 1u << j;
    }
    if (PASS == 0 && valid) need2[cell] = later ? 1 : 0;
    if (jmask == 0) return;
    jmask |= 1u << 8; // last round: nothing new, the rest of the queue is worked off
    // The blocks in lockstep; every lane pops up to LB_POP candidates per round (their snapshot look-ups leave together), the
    // candidates of the warp are collected in shared memory and worked off 32 at a time, so the expensive part (root look-ups,
    // pair trials) always runs on a full warp.
    const unsigned lt = (1u << lane) - 1u;
    uint32_t qn = 0;
    while (jmask) {
        const int j = __ffs(jmask) - 1;
        jmask &= jmask - 1;
        // occ[j], rank[j] by a chain of selects: the arrays stay in registers and the loop body exists once
        unsigned long long oj = occ[0];
        uint32_t rj = rank[0];
#pragma unroll
        for (int k = 1; k < 8; k++) if (j == k) { oj = occ[k]; rj = rank[k]; }
        unsigned long long cand = j < 8 ? (oj & s_tab[loc * 8 + (j & 7)]) : 0ull;
        const uint32_t c0 = cand != 0 ? __ldg(Bp.T.blk_cell0 + rj) : 0u;
        const uint32_t cmj = (uint32_t)((codes >> (6 * (j & 7))) & 63ull) | ((uint32_t)j << 6) | (loc << 15);
        for (;;) {
            const bool more = __any_sync(0xffffffffu, cand != 0);
            if (more) {
                uint32_t B[LB_POP], bit[LB_POP];
                bool keep[LB_POP];
#pragma unroll
                for (int k = 0; k < LB_POP; k++) {
                    keep[k] = cand != 0;
                    bit[k] = keep[k] ? (uint32_t)__ffsll((long long)cand) - 1u : 0u;
                    cand &= cand - 1ull; // 0 stays 0
                    B[k] = c0 + (uint32_t)__popcll(oj & ((1ull << bit[k]) - 1ull));
                }
                if (PASS == 1) {
                    uint32_t sb[LB_POP];
#pragma unroll
                    for (int k = 0; k < LB_POP; k++) sb[k] = keep[k] ? __ldg(snap + B[k]) : myroot;
#pragma unroll
                    for (int k = 0; k < LB_POP; k++) keep[k] = sb[k] != myroot;
                }
#pragma unroll
                for (int k = 0; k < LB_POP; k++) {
                    const unsigned bal = __ballot_sync(0xffffffffu, keep[k]);
                    if (keep[k]) {
                        const uint32_t slot = qn + (uint32_t)__popc(bal & lt);
                        s_qa[w][slot] = cell; s_qb[w][slot] = B[k]; s_qm[w][slot] = cmj | (bit[k] << 9);
                    }
                    qn += (uint32_t)__popc(bal);
                }
                __syncwarp();
            }
            while (qn >= 32 || (!more && j == 8 && qn > 0)) {
                const uint32_t take = qn >= 32 ? 32u : qn;
                qn -= take;
                if (lane < take) link_candidate(P, s_qa[w][qn + lane], s_qb[w][qn + lane], s_qm[w][qn + lane], Q);
                __syncwarp();
            }
            if (!more) break;
        }
    }
}

// one warp per undecided cell pair: the exhaustive test
__global__ void __launch_bounds__(256) k_link_queue(LinkParams P, LinkQueue Q)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t warps_total = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nq = min(*Q.count, Q.capacity); // written by the kernel before this one: no host round trip
    for (uint32_t e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < nq; e += warps_total) {
        const uint32_t cell = Q.a[e], B = Q.b[e], m = Q.meta[e];
        int ix, iy, iz;
        link3::key_cell(P.g, P.cell_key[cell], ix, iy, iz);
        try_link_cells(P, cell, P.cell_start[cell], P.cell_start[cell + 1], ix, iy, iz, B, (int)((m >> 6) & 7) - 2, (int)((m >> 9) & 7) - 2,
                       (int)((m >> 12) & 7) - 2, (int)(m & 63), lane);
    }
}

__global__ void k_flatten(uint32_t *__restrict__ parent, uint32_t ncells, uint32_t *__restrict__ root)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    root[c] = uf_find(parent, c);
}

// per cell: add its particle count / min pid / min original index into its root's records
// cells are in spatial order, so neighbouring lanes often share a root: one atomic triple per (warp, root)
__global__ void k_group_stats(const uint32_t *__restrict__ root, const uint32_t *__restrict__ cell_start, uint32_t ncells,
                              const uint32_t *__restrict__ order, const long long *__restrict__ pid, uint32_t *__restrict__ gsize,
                              unsigned long long *__restrict__ gminpid, uint32_t *__restrict__ gminidx)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = c < ncells;
    unsigned long long mp = ~0ull;
    uint32_t mi = 0xffffffffu, cnt = 0, r = 0xffffffffu;
    if (valid) {
        const uint32_t s = cell_start[c], e = cell_start[c + 1];
        for (uint32_t k = s; k < e; k++) {
            const uint32_t i = order[k];
            const unsigned long long p = pid ? (unsigned long long)pid[i] : (unsigned long long)i;
            mp = p < mp ? p : mp;
            mi = i < mi ? i : mi;
        }
        cnt = e - s;
        r = root[c];
    }
    const unsigned peers = __match_any_sync(0xffffffffu, r);
    const uint32_t tot = __reduce_add_sync(peers, cnt);
    const uint32_t mi_min = __reduce_min_sync(peers, mi);
    const unsigned long long mp_min = warp_min_u64(peers, mp);
    if (valid && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) {
        atomicAdd(&gsize[r], tot);
        atomicMin(&gminpid[r], mp_min);
        atomicMin(&gminidx[r], mi_min);
    }
}

__global__ void k_flag_groups(const uint32_t *__restrict__ root, const uint32_t *__restrict__ gsize, uint32_t ncells, uint32_t minsize,
                              uint32_t *__restrict__ flags)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    flags[c] = (root[c] == c && gsize[c] >= minsize) ? 1u : 0u;
}

// compact the selected roots; build the two sort keys
__global__ void k_compact_groups(const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl, uint32_t ncells,
                                 const uint32_t *__restrict__ gsize, const unsigned long long *__restrict__ gminpid,
                                 uint32_t *__restrict__ g_root, uint64_t *__restrict__ key_pid, uint32_t *__restrict__ val)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells || !flags[c]) return;
    uint32_t g = excl[c];
    g_root[g] = c;
    key_pid[g] = gminpid[c];
    val[g] = g;
}

__global__ void k_size_keys(const uint32_t *__restrict__ val, const uint32_t *__restrict__ g_root, const uint32_t *__restrict__ gsize,
                            uint32_t ng, uint32_t maxsize, uint64_t *__restrict__ key_size)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ng) return;
    key_size[k] = (uint64_t)(maxsize - gsize[g_root[val[k]]]);
}

__global__ void k_assign_gids(const uint32_t *__restrict__ val, const uint32_t *__restrict__ g_root, uint32_t ng, uint32_t *__restrict__ gid_of_cell)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ng) return;
    gid_of_cell[g_root[val[k]]] = k + 1;
}

__global__ void k_write_pfof(const uint32_t *__restrict__ order, const uint32_t *__restrict__ cell_of, const uint32_t *__restrict__ root,
                             const uint32_t *__restrict__ gid_of_cell, size_t n, long long *__restrict__ pfof, unsigned long long *__restrict__ ngrouped)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t g = 0;
    if (k < n) {
        g = gid_of_cell[root[cell_of[k]]];
        pfof[order[k]] = (long long)g;
    }
    unsigned m = __ballot_sync(0xffffffffu, g != 0);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(ngrouped, (unsigned long long)__popc(m));
}

__global__ void k_reduce_max_u32(const uint32_t *__restrict__ a, uint32_t n, uint32_t *__restrict__ out)
{
    uint32_t m = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) m = max(m, a[i]);
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}
__global__ void k_reduce_max_u64(const unsigned long long *__restrict__ a, const uint32_t *__restrict__ flags, uint32_t n,
                                 unsigned long long *__restrict__ out)
{
    unsigned long long m = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        if (flags[i]) m = a[i] > m ? a[i] : m;
    for (int o = 16; o; o >>= 1) {
        unsigned long long y = __shfl_xor_sync(0xffffffffu, m, o);
        m = y > m ? y : m;
    }
    if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

static int bits_for(uint64_t v)
{
    int b = 0;
    while (v) { b++; v >>= 1; }
    return b < 1 ? 1 : b;
}

static void init_offsets()
{
    static bool done = false;
    if (done) return;
    signed char h[62][3];
    int k = 0;
    for (int pass = 0; pass < 2; pass++)
        for (int dz = -2; dz <= 2; dz++)
            for (int dy = -2; dy <= 2; dy++)
                for (int dx = -2; dx <= 2; dx++) {
                    bool half = dz > 0 || (dz == 0 && dy > 0) || (dz == 0 && dy == 0 && dx > 0);
                    if (!half) continue;
                    int m = std::max(std::abs(dx), std::max(std::abs(dy), std::abs(dz)));
                    if ((pass == 0) != (m <= 1)) continue;
                    h[k][0] = (signed char)dx; h[k][1] = (signed char)dy; h[k][2] = (signed char)dz;
                    k++;
                }
    VR_REQUIRE(k == 62, VRGPU_EINVAL, "offset table");
    VR_CUDA(cudaMemcpyToSymbol(c_offsets, h, sizeof(h)));
    done = true;
}

// Assign ids 1..G to the selected roots ordered by (size desc, min pid asc); returns G.
// gid_of_cell[ncells] receives the id of each ROOT cell (0 elsewhere).
uint32_t order_groups(vrgpu_ctx *ctx, const uint32_t *root, const uint32_t *gsize, const unsigned long long *gminpid,
                      uint32_t ncells, uint32_t minsize, uint32_t *gid_of_cell)
{
    const int T = 256;
    DBuf<uint32_t> flags(ctx, ncells), excl(ctx, ncells), total(ctx, 4);
    VR_LAUNCH(ctx, k_flag_groups, grid_for(ncells, T), T, 0, root, gsize, ncells, minsize, flags.p);
    exclusive_scan_u32(ctx, flags.p, excl.p, ncells, total.p);
    DBuf<unsigned long long> maxpid(ctx, 1);
    VR_CUDA(cudaMemsetAsync(total.p + 1, 0, sizeof(uint32_t), ctx->stream));
    VR_CUDA(cudaMemsetAsync(maxpid.p, 0, sizeof(unsigned long long), ctx->stream));
    VR_LAUNCH(ctx, k_reduce_max_u32, 296, T, 0, gsize, ncells, total.p + 1);
    VR_LAUNCH(ctx, k_reduce_max_u64, 296, T, 0, gminpid, flags.p, ncells, maxpid.p);
    uint32_t h[2];
    unsigned long long hmaxpid;
    VR_CUDA(cudaMemcpyAsync(h, total.p, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(&hmaxpid, maxpid.p, sizeof(hmaxpid), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint32_t ng = h[0], maxsize = h[1];
    VR_CUDA(cudaMemsetAsync(gid_of_cell, 0, sizeof(uint32_t) * ncells, ctx->stream));
    if (ng == 0) return 0;
    DBuf<uint32_t> g_root(ctx, ng), val(ctx, ng), val_tmp(ctx, ng);
    DBuf<uint64_t> key(ctx, ng), key_tmp(ctx, ng);
    VR_LAUNCH(ctx, k_compact_groups, grid_for(ncells, T), T, 0, flags.p, excl.p, ncells, gsize, gminpid, g_root.p, key.p, val.p);
    radix_sort_pairs(ctx, key.p, val.p, key_tmp.p, val_tmp.p, ng, bits_for(hmaxpid));
    VR_LAUNCH(ctx, k_size_keys, grid_for(ng, T), T, 0, val.p, g_root.p, gsize, ng, maxsize, key.p);
    radix_sort_pairs(ctx, key.p, val.p, key_tmp.p, val_tmp.p, ng, bits_for(maxsize));
    VR_LAUNCH(ctx, k_assign_gids, grid_for(ng, T), T, 0, val.p, g_root.p, ng, gid_of_cell);
    return ng;
}

void fof3d_device(vrgpu_ctx *ctx, const DeviceParticles &P, double b, double period, long long minsize, Fof3dState &S)
{
    fof3d_link(ctx, P, b, period, nullptr, S);
    fof3d_label(ctx, P, minsize, S);
}

// cells + union-find: fills S.order / cell_of / cell_start / cell_key / root (root cell of every non-empty cell)
void fof3d_link(vrgpu_ctx *ctx, const DeviceParticles &P, double b, double period, const int *periodic_axes, Fof3dState &S)
{
    const size_t n = P.n;
    const int T = 256;
    S.n = n;
    S.ngroups = 0;
    S.ngrouped = 0;
    S.ncells = 0;
    if (n == 0) return;
    VR_REQUIRE(n < (size_t)0xfffffff0u, VRGPU_EINVAL, "more than 2^32 particles on one device");
    VR_REQUIRE(b > 0, VRGPU_EINVAL, "linking length must be positive");
    init_offsets();

    // ---- grid ----
    Grid g{};
    g.period = period > 0 ? period : 0.0;
    bool any_open = false;
    for (int k = 0; k < 3; k++) {
        g.periodic[k] = g.period > 0 && (!periodic_axes || periodic_axes[k]);
        any_open = any_open || !g.periodic[k];
    }
    const double cmax = b / std::sqrt(3.0) * (1.0 - 1e-8);
    double lohi[6] = {0, 0, 0, 0, 0, 0};
    if (any_open) {
        DBuf<double> mm(ctx, 6);
        unsigned long long init[6] = {~0ull, ~0ull, ~0ull, 0ull, 0ull, 0ull};
        VR_CUDA(cudaMemcpyAsync(mm.p, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
        VR_LAUNCH(ctx, k_minmax, std::min<unsigned>(grid_for(n, T), 1184u), T, 0, P.pos.p, n, mm.p);
        unsigned long long enc[6];
        VR_CUDA(cudaMemcpyAsync(enc, mm.p, sizeof(enc), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < 6; k++) {
            unsigned long long u = enc[k];
            u = (u & 0x8000000000000000ull) ? (u & 0x7fffffffffffffffull) : ~u;
            std::memcpy(&lohi[k], &u, 8);
        }
    }
    // Periodic axes: whole 4-cell blocks, so that the grid wraps block on block -- as long as cells three apart still cannot link
    // (2 c >= b).  Boxes below ~15 linking lengths may have no such cell count: they keep the plain count and take the hash path.
    bool use_blocks = true;
    double mper = 0;
    if (g.period > 0) {
        mper = std::ceil(g.period / cmax);
        VR_REQUIRE(mper < 2.0e6, VRGPU_EUNSUPPORTED, "linking length too small for the cell key (period/b > ~1e6)");
        VR_REQUIRE(mper >= 8, VRGPU_EUNSUPPORTED, "periodic box smaller than ~5 linking lengths is not supported");
        const double m4 = std::ceil(mper / 4.0) * 4.0;
        if (2.0 * (g.period / m4) >= b * (1.0 + 1e-6)) mper = m4; else use_blocks = false;
    }
    for (int k = 0; k < 3; k++) {
        if (g.periodic[k]) {
            g.lo[k] = 0.0;
            g.nc[k] = (int)mper;
            g.c[k] = g.period / g.nc[k];
            g.inv_c[k] = (double)g.nc[k] / g.period;
        } else {
            g.lo[k] = lohi[k];
            double ext = lohi[3 + k] - lohi[k];
            VR_REQUIRE(std::isfinite(ext), VRGPU_EINVAL, "non-finite positions");
            double m = std::floor(ext / cmax) + 1.0;
            VR_REQUIRE(m < 2.0e6, VRGPU_EUNSUPPORTED, "linking length too small for the cell key (extent/b > ~1e6)");
            g.nc[k] = (int)m;
            g.c[k] = cmax;
            g.inv_c[k] = 1.0 / cmax;
        }
    }
    link3::Geom geom{};
    geom.ncx = g.nc[0]; geom.ncy = g.nc[1]; geom.ncz = g.nc[2];
    geom.nbx = (g.nc[0] + 3) / 4; geom.nby = (g.nc[1] + 3) / 4; geom.nbz = (g.nc[2] + 3) / 4;
    geom.perx = g.periodic[0]; geom.pery = g.periodic[1]; geom.perz = g.periodic[2];
    const uint64_t nblk_total = (uint64_t)geom.nbx * (uint64_t)geom.nby * (uint64_t)geom.nbz;
    const int nbits = bits_for(nblk_total - 1) + 6;
    // the block bitmap and its rank take 3 bits per block of the grid: up to 2^33 blocks (3 GiB)
    if (nblk_total > (1ull << 33)) use_blocks = false;
    const size_t nwords = (size_t)(nblk_total / 64 + 1);
    S.grid = g;

    // ---- keys + sort ----
    DBuf<uint64_t> keys(ctx, n), keys_tmp(ctx, n);
    DBuf<uint32_t> order_tmp(ctx, n);
    S.order.alloc(ctx, n);
    DBuf<unsigned int> nbad(ctx, 1);
    VR_CUDA(cudaMemsetAsync(nbad.p, 0, sizeof(unsigned int), ctx->stream));
    DBuf<uint32_t> blk_cell0, prefix;
    DBuf<unsigned long long> cell_rec, blk_occ, bitmap;
    uint32_t nblocks = 0;
    {
        StageTimer t(ctx);
        { KScope k(ctx, "fof3d.cell_keys"); VR_LAUNCH(ctx, k_cell_keys, grid_for(n, T), T, 0, P.pos.p, n, g, geom, keys.p, S.order.p, nbad.p); }
        { KScope k(ctx, "fof3d.radix_sort"); radix_sort_pairs(ctx, keys.p, S.order.p, keys_tmp.p, order_tmp.p, n, nbits); }
        keys_tmp.reset(); order_tmp.reset();
        S.sx.alloc(ctx, n); S.sy.alloc(ctx, n); S.sz.alloc(ctx, n);
        { KScope k(ctx, "fof3d.gather_pos"); VR_LAUNCH(ctx, k_gather_pos, grid_for(n, T), T, 0, P.pos.p, S.order.p, n, S.sx.p, S.sy.p, S.sz.p); }
        // cells and blocks: one 64-bit scan numbers both
        DBuf<unsigned long long> heads(ctx, n), total(ctx, 1);
        { KScope k(ctx, "fof3d.cell_table"); VR_LAUNCH(ctx, k_flag_heads, grid_for(n, T), T, 0, keys.p, n, heads.p);
        exclusive_scan_u64(ctx, heads.p, heads.p, n, total.p); }
        unsigned long long htot = 0;
        unsigned int hbad = 0;
        VR_CUDA(cudaMemcpyAsync(&htot, total.p, sizeof(htot), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(&hbad, nbad.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        VR_REQUIRE(hbad == 0, VRGPU_EINVAL, "positions are NaN or outside [0, period]");
        const uint32_t hn = (uint32_t)(htot & 0xffffffffull), hb = (uint32_t)(htot >> 32);
        S.ncells = hn;
        nblocks = hb;
        S.cell_of.alloc(ctx, n);
        S.cell_start.alloc(ctx, (size_t)hn + 1);
        S.cell_key.alloc(ctx, hn);
        BlockArrays BA{};
        KScope k(ctx, "fof3d.cell_table");
        if (use_blocks) {
            cell_rec.alloc(ctx, hn); blk_cell0.alloc(ctx, hb); blk_occ.alloc(ctx, hb);
            bitmap.alloc(ctx, nwords); prefix.alloc(ctx, nwords);
            VR_CUDA(cudaMemsetAsync(blk_occ.p, 0, sizeof(unsigned long long) * hb, ctx->stream));
            VR_CUDA(cudaMemsetAsync(bitmap.p, 0, sizeof(unsigned long long) * nwords, ctx->stream));
            BA = BlockArrays{cell_rec.p, blk_cell0.p, blk_occ.p, bitmap.p};
        }
        VR_LAUNCH(ctx, k_fill_cells, grid_for(n, T), T, 0, keys.p, heads.p, n, geom, S.cell_of.p, S.cell_start.p, S.cell_key.p, hn, BA);
        if (use_blocks) {
            VR_LAUNCH(ctx, k_bitmap_counts, grid_for(nwords, T), T, 0, bitmap.p, nwords, prefix.p);
            exclusive_scan_u32(ctx, prefix.p, prefix.p, nwords, nullptr);
        }
        ctx->stats.sort_ms += t.stop();
    }
    keys.reset();
    const uint32_t ncells = S.ncells;
    ctx->stats.n_cells = ncells;

    // ---- link ----
    DBuf<uint32_t> parent(ctx, ncells);
    {
        StageTimer t(ctx);
        size_t h = 1;
        if (!use_blocks) while (h < 2 * (size_t)ncells) h <<= 1;
        DBuf<uint64_t> hkeys(ctx, h);
        DBuf<uint32_t> hvals(ctx, h);
        if (use_blocks) {
            VR_LAUNCH(ctx, k_uf_init, grid_for(ncells, T), T, 0, parent.p, ncells);
        } else {
            KScope k(ctx, "fof3d.hash_build");
            VR_LAUNCH(ctx, k_hash_fill, grid_for(h, T), T, 0, hkeys.p, h);
            VR_LAUNCH(ctx, k_hash_insert, grid_for(ncells, T), T, 0, S.cell_key.p, ncells, hkeys.p, hvals.p, (uint32_t)(h - 1), parent.p);
        }
        LinkParams L{};
        L.sx = S.sx.p; L.sy = S.sy.p; L.sz = S.sz.p;
        L.cell_start = S.cell_start.p; L.cell_key = S.cell_key.p; L.ncells = ncells;
        L.parent = parent.p; L.hkeys = hkeys.p; L.hvals = hvals.p; L.hmask = (uint32_t)(h - 1);
        L.ncx = g.nc[0]; L.ncy = g.nc[1]; L.ncz = g.nc[2];
        L.cx = g.c[0]; L.cy = g.c[1]; L.cz = g.c[2];
        L.lox = g.lo[0]; L.loy = g.lo[1]; L.loz = g.lo[2];
        L.period = g.period; L.perx = g.periodic[0]; L.pery = g.periodic[1]; L.perz = g.periodic[2];
        L.b2 = b * b; // KDFOF.cxx:14
        L.g = geom;
        if (use_blocks) {
            // candidate table: built once per context on the host (pure integer function, link3.cuh)
            static std::vector<unsigned long long> host_tab = [] {
                std::vector<uint64_t> t(link3::TABLE_WORDS);
                link3::build_table(t.data());
                return std::vector<unsigned long long>(t.begin(), t.end());
            }();
            DBuf<unsigned long long> tab(ctx, link3::TABLE_WORDS);
            VR_CUDA(cudaMemcpyAsync(tab.p, host_tab.data(), sizeof(unsigned long long) * link3::TABLE_WORDS, cudaMemcpyHostToDevice, ctx->stream));
            const uint32_t qcap = (uint32_t)std::min<size_t>((size_t)8 * ncells + 1024, (size_t)0x7fffffff);
            DBuf<uint32_t> qa(ctx, qcap), qb(ctx, qcap), qm(ctx, qcap);
            DBuf<unsigned int> qcount(ctx, 1);
            LinkQueue Q{qa.p, qb.p, qm.p, qcount.p, qcap};
            BlockParams Bp{link3::Tables{bitmap.p, prefix.p, blk_occ.p, blk_cell0.p}, cell_rec.p, tab.p};
            KScope k(ctx, "fof3d.link_cells");
            VR_CUDA(cudaMemsetAsync(qcount.p, 0, sizeof(unsigned int), ctx->stream));
            S.root.alloc(ctx, ncells);
            DBuf<uint32_t> blk_root(ctx, nblocks);
            DBuf<uint8_t> need2(ctx, ncells);
            VR_LAUNCH(ctx, k_link_blocks<0>, grid_for(ncells, LB_THREADS), LB_THREADS, 0, L, Bp, (const uint32_t *)nullptr, (const uint32_t *)nullptr, need2.p, Q);
            VR_LAUNCH(ctx, k_flatten, grid_for(ncells, T), T, 0, parent.p, ncells, S.root.p); // snapshot of the roots after pass 0
            VR_LAUNCH(ctx, k_block_roots, grid_for(nblocks, T), T, 0, S.root.p, blk_cell0.p, nblocks, ncells, blk_root.p);
            VR_LAUNCH(ctx, k_link_blocks<1>, grid_for(ncells, LB_THREADS), LB_THREADS, 0, L, Bp, (const uint32_t *)S.root.p, (const uint32_t *)blk_root.p, need2.p, Q);
            // dense cell pairs the short trial left undecided: one warp per pair, the count is read on the device
            VR_LAUNCH(ctx, k_link_queue, (unsigned)ctx->sm_count * 16u, 256, 0, L, Q);
            unsigned int nq = 0;
            VR_CUDA(cudaMemcpyAsync(&nq, qcount.p, sizeof(nq), cudaMemcpyDeviceToHost, ctx->stream));
            { KScope kf(ctx, "fof3d.flatten"); VR_LAUNCH(ctx, k_flatten, grid_for(ncells, T), T, 0, parent.p, ncells, S.root.p); }
            ctx->stats.link_ms += t.stop(); // synchronises: nq has arrived
            ctx->stats.n_link_queue += std::min(nq, qcap);
        } else {
            // one warp per cell, grid a multiple of the SM count
            const unsigned blocks = std::max(1u, std::min<unsigned>((ncells + 7) / 8, (unsigned)ctx->sm_count * 32u));
            { KScope k(ctx, "fof3d.link_cells");
              VR_LAUNCH(ctx, k_link_cells, blocks, 256, 0, L, 0, 13);
              VR_LAUNCH(ctx, k_link_cells, blocks, 256, 0, L, 13, 62); }
            S.root.alloc(ctx, ncells);
            { KScope k(ctx, "fof3d.flatten"); VR_LAUNCH(ctx, k_flatten, grid_for(ncells, T), T, 0, parent.p, ncells, S.root.p); }
            ctx->stats.link_ms += t.stop();
        }
    }
}

// group sizes / min pid per root cell, drop < minsize, ids by (size desc, min pid asc), pfof in caller order
void fof3d_label(vrgpu_ctx *ctx, const DeviceParticles &P, long long minsize, Fof3dState &S)
{
    const size_t n = P.n;
    const int T = 256;
    S.pfof.alloc(ctx, n ? n : 1);
    if (n == 0) return;
    const uint32_t ncells = S.ncells;
    {
        StageTimer t(ctx);
        DBuf<uint32_t> gsize(ctx, ncells), gminidx(ctx, ncells), gid_of_cell(ctx, ncells);
        DBuf<unsigned long long> gminpid(ctx, ncells), ngrouped(ctx, 1);
        VR_CUDA(cudaMemsetAsync(gsize.p, 0, sizeof(uint32_t) * ncells, ctx->stream));
        VR_CUDA(cudaMemsetAsync(gminidx.p, 0xff, sizeof(uint32_t) * ncells, ctx->stream));
        VR_CUDA(cudaMemsetAsync(gminpid.p, 0xff, sizeof(unsigned long long) * ncells, ctx->stream));
        VR_CUDA(cudaMemsetAsync(ngrouped.p, 0, sizeof(unsigned long long), ctx->stream));
        { KScope k(ctx, "fof3d.group_stats"); VR_LAUNCH(ctx, k_group_stats, grid_for(ncells, T), T, 0, S.root.p, S.cell_start.p, ncells, S.order.p,
                  P.has_pid ? P.pid.p : (const long long *)nullptr, gsize.p, gminpid.p, gminidx.p); }
        uint32_t ms = minsize < 1 ? 1u : (minsize > 0xffffffffLL ? 0xffffffffu : (uint32_t)minsize);
        uint32_t ng;
        { KScope k(ctx, "fof3d.order_groups"); ng = order_groups(ctx, S.root.p, gsize.p, gminpid.p, ncells, ms, gid_of_cell.p); }
        KScope kw(ctx, "fof3d.write_pfof");
        VR_LAUNCH(ctx, k_write_pfof, grid_for(n, T), T, 0, S.order.p, S.cell_of.p, S.root.p, gid_of_cell.p, n, S.pfof.p, ngrouped.p);
        unsigned long long hg = 0;
        VR_CUDA(cudaMemcpyAsync(&hg, ngrouped.p, sizeof(hg), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        S.ngroups = ng;
        S.ngrouped = (long long)hg;
        ctx->stats.n_groups3d = ng;
        ctx->stats.n_grouped = (long long)hg;
        ctx->stats.label_ms += t.stop();
    }
}

// ---------------------------------------------------------------------------------------------------
// AdjustStructureForPeriod (src/search.cxx:856-895)
// ---------------------------------------------------------------------------------------------------
// first member (lowest index) of every group; the big groups (low ids) are reduced per CTA in shared memory first
constexpr int GF_IDS = 4096, GF_THREADS = 512;
__global__ void __launch_bounds__(GF_THREADS) k_group_first(const long long *__restrict__ pfof, size_t n, uint32_t *__restrict__ first)
{
    __shared__ uint32_t sf[GF_IDS];
    for (int k = threadIdx.x; k < GF_IDS; k += GF_THREADS) sf[k] = 0xffffffffu;
    __syncthreads();
    const size_t per = (n + gridDim.x - 1) / gridDim.x;
    const size_t i0 = (size_t)blockIdx.x * per, i1 = min(n, i0 + per);
    for (size_t i = i0 + threadIdx.x; i < i1; i += GF_THREADS) {
        const long long g = pfof[i];
        if (g > 0) {
            if (g <= GF_IDS) atomicMin(&sf[g - 1], (uint32_t)i);
            else atomicMin(&first[g], (uint32_t)i);
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < GF_IDS; k += GF_THREADS)
        if (sf[k] != 0xffffffffu) atomicMin(&first[k + 1], sf[k]);
}

__global__ void k_adjust_period(double *__restrict__ pos, const long long *__restrict__ pfof, size_t n, const uint32_t *__restrict__ first,
                                double period, uint32_t *__restrict__ moved_idx, double *__restrict__ moved_xyz, unsigned int *__restrict__ nmoved,
                                int pass)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long g = pfof[i];
    if (g == 0) return;
    uint32_t r = first[g];
    if (r == i) return;
    // pass 0 counts and computes, pass 1 is not needed: the reference particle never moves, so in-place is race free
    double np[3];
    bool moved = false;
    for (int j = 0; j < 3; j++) {
        double x = pos[3 * i + j];
        double diff = __dsub_rn(pos[3 * (size_t)r + j], x);
        if (diff < -0.5 * period) { x = __dsub_rn(x, period); moved = true; }
        else if (diff > 0.5 * period) { x = __dadd_rn(x, period); moved = true; }
        np[j] = x;
    }
    if (moved) {
        unsigned int slot = atomicAdd(nmoved, 1u);
        moved_idx[slot] = (uint32_t)i;
        for (int j = 0; j < 3; j++) { moved_xyz[3 * (size_t)slot + j] = np[j]; pos[3 * i + j] = np[j]; }
    }
    (void)pass;
}

__global__ void k_group_minpid(const long long *__restrict__ pfof, const long long *__restrict__ pid, size_t n, unsigned long long *__restrict__ minpid)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const long long g = i < n ? pfof[i] : 0ll;
    const bool on = g > 0;
    const unsigned peers = __match_any_sync(0xffffffffu, on ? (unsigned long long)g : 0ull); // one atomic per (warp, group)
    const unsigned long long mn = warp_min_u64(peers, on ? (unsigned long long)pid[i] : ~0ull);
    if (on && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) atomicMin(&minpid[g], mn);
}
__global__ void k_group_first_by_pid(const long long *__restrict__ pfof, const long long *__restrict__ pid, size_t n,
                                     const unsigned long long *__restrict__ minpid, uint32_t *__restrict__ first)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long g = pfof[i];
    if (g > 0 && (unsigned long long)pid[i] == minpid[g]) atomicMin(&first[g], (uint32_t)i);
}

size_t adjust_for_period_device(vrgpu_ctx *ctx, DeviceParticles &P, const long long *pfof_dev, long long ngroups, double period,
                                DBuf<uint32_t> &moved_idx, DBuf<double> &moved_xyz, bool by_pid)
{
    if (P.n == 0 || ngroups <= 0 || !(period > 0)) return 0;
    const int T = 256;
    DBuf<uint32_t> first(ctx, (size_t)ngroups + 1);
    VR_CUDA(cudaMemsetAsync(first.p, 0xff, sizeof(uint32_t) * ((size_t)ngroups + 1), ctx->stream));
    if (by_pid && P.has_pid) {
        // multi-GPU runs: the array order of a consolidated group is not the caller's, so the reference particle is the
        // member with the smallest pid (== the first member whenever pids follow the input order)
        DBuf<unsigned long long> minpid(ctx, (size_t)ngroups + 1);
        VR_CUDA(cudaMemsetAsync(minpid.p, 0xff, sizeof(unsigned long long) * ((size_t)ngroups + 1), ctx->stream));
        VR_LAUNCH(ctx, k_group_minpid, grid_for(P.n, T), T, 0, pfof_dev, P.pid.p, P.n, minpid.p);
        VR_LAUNCH(ctx, k_group_first_by_pid, grid_for(P.n, T), T, 0, pfof_dev, P.pid.p, P.n, minpid.p, first.p);
    } else
        VR_LAUNCH(ctx, k_group_first, (unsigned)std::max<size_t>(1, std::min<size_t>((size_t)ctx->sm_count * 4, (P.n + GF_THREADS - 1) / GF_THREADS)), GF_THREADS,
                  0, pfof_dev, P.n, first.p);
    // worst case every grouped particle moves; allocate for n (lists are compacted by the atomic counter)
    moved_idx.alloc(ctx, P.n);
    moved_xyz.alloc(ctx, 3 * P.n);
    DBuf<unsigned int> nmoved(ctx, 1);
    VR_CUDA(cudaMemsetAsync(nmoved.p, 0, sizeof(unsigned int), ctx->stream));
    VR_LAUNCH(ctx, k_adjust_period, grid_for(P.n, T), T, 0, P.pos.p, pfof_dev, P.n, first.p, period, moved_idx.p, moved_xyz.p, nmoved.p, 0);
    unsigned int h = 0;
    VR_CUDA(cudaMemcpyAsync(&h, nmoved.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return h;
}

} // namespace vrgpu
