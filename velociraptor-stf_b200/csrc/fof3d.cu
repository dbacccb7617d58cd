// velociraptor-stf_b200/csrc/fof3d.cu -- 3D friends-of-friends on the device.
//
// Replaces NBody::KDTree::FOF (NBodylib/src/KDTree/KDFOF.cxx:12-129) + the leaf/split-node ball searches
// (KDLeafNode.cxx:518-573, KDSplitNode.cxx:865-891, 1496-1540) as called by SearchFullSet (src/search.cxx:203).
//
// FOF groups are the transitive closure of the pair predicate, so the kd-tree BFS is not reproduced; instead
//   1. particles are binned into FINE cells of edge c < b/sqrt(3): two particles in one cell are always linked,
//      so the union-find runs over non-empty CELLS, not particles;
//   2. particles are radix-sorted by cell key (sort.cu), positions gathered into cell order (SoA, coalesced);
//   3. a hash table maps cell key -> cell id; one warp per cell probes its 62 half-space neighbours (|d|<=2),
//      skips neighbours already in the same set, otherwise tests particle pairs with the reference's exact
//      fp64 predicate until one link is found and unites the two cells (lock-free atomicCAS union, path halving);
//   4. roots are flattened, group sizes / min pid accumulated per root, groups >= minsize ordered by
//      (size desc, min pid asc) and pfof written in the caller's particle order.
// Pair predicate: d2 = dx*dx + dy*dy + dz*dz summed left to right in fp64 WITHOUT fma, link iff d2 < b*b
// (KDLeafNode.cxx:556-558, DistFunc.h:27-32, KDFOF.cxx:14).  Periodic images: the target is shifted by +-period
// before the subtraction (KDSplitNode.cxx:1500-1539); either particle of a pair can be the BFS target in the
// reference, so both orientations are evaluated.
#include "fof.cuh"
#include "groups.cuh"
#include "ufhash.cuh"

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
};

__constant__ signed char c_offsets[62][3];

// ---------------------------------------------------------------------------------------------------
// binning
// ---------------------------------------------------------------------------------------------------
__global__ void k_cell_keys(const double *__restrict__ pos, size_t n, Grid g, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals,
                            unsigned int *__restrict__ nbad)
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
    keys[i] = ((uint64_t)iz * (uint64_t)g.nc[1] + (uint64_t)iy) * (uint64_t)g.nc[0] + (uint64_t)ix;
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

__global__ void k_flag_heads(const uint64_t *__restrict__ keys, size_t n, uint32_t *__restrict__ flags)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    flags[k] = (k == 0 || keys[k] != keys[k - 1]) ? 1u : 0u;
}

__global__ void k_fill_cells(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl,
                             size_t n, uint32_t *__restrict__ cell_of, uint32_t *__restrict__ cell_start, uint64_t *__restrict__ cell_key,
                             uint32_t ncells)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint32_t f = flags[k];
    uint32_t c = excl[k] - (f ? 0u : 1u);
    cell_of[k] = c;
    if (f) {
        cell_start[c] = (uint32_t)k;
        cell_key[c] = keys[k];
    }
    if (k == n - 1) cell_start[ncells] = (uint32_t)n;
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

// One warp per cell.  off_begin/off_end select the slice of the 62 half-space offsets handled by this launch:
// the 13 face/edge/corner neighbours first, the 49 second-shell ones afterwards (by then most are already united).
__global__ void __launch_bounds__(256) k_link_cells(LinkParams P, int off_begin, int off_end)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t warps_total = (gridDim.x * blockDim.x) >> 5;
    const uint32_t warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t nxy = (uint64_t)P.ncx * (uint64_t)P.ncy;
    for (uint32_t cell = warp_global; cell < P.ncells; cell += warps_total) {
        const uint64_t key = P.cell_key[cell];
        const int iz = (int)(key / nxy);
        const uint64_t rem = key - (uint64_t)iz * nxy;
        const int iy = (int)(rem / (uint64_t)P.ncx);
        const int ix = (int)(rem - (uint64_t)iy * (uint64_t)P.ncx);
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
                    uint64_t nkey = ((uint64_t)nz * (uint64_t)P.ncy + (uint64_t)ny) * (uint64_t)P.ncx + (uint64_t)nx;
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
// row-table linker.  Cells are sorted by key = (iz*ncy + iy)*ncx + ix, so the cells of one (iy,iz) ROW are contiguous
// and ordered by ix.  A dense table row -> [first,last) cell (4 B x 2 per row; 39 MB at 256^3, L2 resident) turns the 62
// neighbour-cell hash probes of a cell into 12 row look-ups + short scans (the 5 x-neighbours of a row are adjacent).
// ---------------------------------------------------------------------------------------------------
constexpr uint32_t ROW_NONE = 0xffffffffu;

__global__ void k_row_table(const uint64_t *__restrict__ cell_key, uint32_t ncells, uint64_t ncx, uint32_t *__restrict__ row_first,
                            uint32_t *__restrict__ row_last)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    const uint64_t r = cell_key[c] / ncx;
    if (c == 0 || cell_key[c - 1] / ncx != r) row_first[r] = c;
    if (c == ncells - 1 || cell_key[c + 1] / ncx != r) row_last[r] = c + 1;
}

// reach = 1: only the 13 adjacent half-space cells (cheap first pass that unites most of a halo), reach = 2: all 62.
__global__ void __launch_bounds__(256) k_link_rows(LinkParams P, const uint32_t *__restrict__ row_first, const uint32_t *__restrict__ row_last,
                                                   int reach)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t warps_total = (gridDim.x * blockDim.x) >> 5;
    const uint32_t warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t ncx = (uint64_t)P.ncx;
    for (uint32_t cell = warp_global; cell < P.ncells; cell += warps_total) {
        const uint64_t key = P.cell_key[cell];
        const uint64_t row = key / ncx;
        const int ix = (int)(key - row * ncx);
        const int iz = (int)(row / (uint64_t)P.ncy);
        const int iy = (int)(row - (uint64_t)iz * (uint64_t)P.ncy);
        const uint32_t a0 = P.cell_start[cell], a1 = P.cell_start[cell + 1];
        // lane -> (role, segment): role 0..11 = neighbour rows of the half space, role 12 = own row (dx > 0);
        // segment 0 = the part of the x-interval inside [0,ncx), segment 1 = the part that wraps (periodic x only)
        const int role = (int)(lane & 15), seg = (int)(lane >> 4);
        int dy = 0, dz = 0;
        bool ok = role <= 12;
        if (role < 5) { dz = 1; dy = role - 2; }
        else if (role < 10) { dz = 2; dy = role - 7; }
        else if (role < 12) { dz = 0; dy = role - 9; }
        if (ok && (abs(dy) > reach || dz > reach)) ok = false;
        int xlo = role == 12 ? ix + 1 : ix - reach, xhi = ix + reach; // unwrapped x-interval
        int xoff = 0, sx = 0, sy = 0, sz = 0;
        if (ok) {
            if (seg == 0) { xlo = max(xlo, 0); xhi = min(xhi, P.ncx - 1); }
            else if (!P.perx) ok = false;
            else if (xlo < 0) { xhi = min(xhi, -1) + P.ncx; xlo += P.ncx; xoff = -P.ncx; sx = 1; }
            else if (xhi >= P.ncx) { xlo = max(xlo, P.ncx) - P.ncx; xhi -= P.ncx; xoff = P.ncx; sx = -1; }
            else ok = false;
            if (xlo > xhi) ok = false;
        }
        int ny = iy + dy, nz = iz + dz;
        if (ok) {
            if (ny < 0) { if (P.pery) { ny += P.ncy; sy = 1; } else ok = false; }
            else if (ny >= P.ncy) { if (P.pery) { ny -= P.ncy; sy = -1; } else ok = false; }
            if (nz >= P.ncz) { if (P.perz) { nz -= P.ncz; sz = -1; } else ok = false; }
        }
        uint32_t c0 = 0, c1 = 0;
        uint64_t rowbase = 0;
        if (ok) {
            const uint64_t r = (uint64_t)nz * (uint64_t)P.ncy + (uint64_t)ny;
            rowbase = r * ncx;
            uint32_t f = row_first[r];
            if (f != ROW_NONE) {
                uint32_t l = row_last[r];
                const uint64_t klo = rowbase + (uint64_t)xlo, khi = rowbase + (uint64_t)xhi;
                if (role == 12 && seg == 0) f = cell + 1; // own row, forward: the next cells in the array
                else {
                    // lower bound of klo in cell_key[f, l)
                    uint32_t lo = f, hi = l;
                    while (lo < hi) {
                        uint32_t mid = (lo + hi) >> 1;
                        if (__ldg(&P.cell_key[mid]) < klo) lo = mid + 1; else hi = mid;
                    }
                    f = lo;
                }
                uint32_t e = f;
                while (e < l && __ldg(&P.cell_key[e]) <= khi) e++;
                c0 = f; c1 = e;
            }
        }
        const int scode = (sx + 1) | ((sy + 1) << 2) | ((sz + 1) << 4);
        unsigned has = __ballot_sync(0xffffffffu, c0 < c1);
        while (has) {
            const int src = __ffs(has) - 1;
            const uint32_t B = __shfl_sync(0xffffffffu, c0, src);
            const int sc = __shfl_sync(0xffffffffu, scode, src);
            const int ddy = __shfl_sync(0xffffffffu, dy, src), ddz = __shfl_sync(0xffffffffu, dz, src);
            const int xo = __shfl_sync(0xffffffffu, xoff, src);
            const unsigned long long rb = __shfl_sync(0xffffffffu, (unsigned long long)rowbase, src);
            const int ddx = (int)(P.cell_key[B] - rb) + xo - ix;
            if ((int)lane == src) c0++;
            try_link_cells(P, cell, a0, a1, ix, iy, iz, B, ddx, ddy, ddz, sc, lane);
            has = __ballot_sync(0xffffffffu, c0 < c1);
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// thread-per-cell linker.  Most non-empty cells of a cosmological box hold ONE particle and have no neighbour at all, so a
// warp per cell spends hundreds of instructions on bookkeeping for nothing (ncu, profiles/r01b: 5.8e9 warp instructions,
// 1 % of them fp64).  Here every THREAD owns a cell: it walks its 12 neighbour rows + its own row through the row table,
// and for every neighbour cell in a different set it tests up to 16 pairs inline (enough for sparse cells and for most
// dense neighbours, which link at the first try); only cell pairs that are still undecided go to a queue that a
// warp-per-pair kernel finishes with the exhaustive test.
// ---------------------------------------------------------------------------------------------------
struct LinkQueue {
    uint32_t *a, *b, *meta;     // meta: sc (6 bits) | (ddx+2)<<6 | (ddy+2)<<9 | (ddz+2)<<12
    unsigned int *count;
    uint32_t capacity;
};

__global__ void k_cell_coords(const uint64_t *__restrict__ cell_key, uint32_t ncells, uint64_t ncx, uint32_t *__restrict__ cell_row,
                              int *__restrict__ cell_ix, uint32_t *__restrict__ row_first, uint32_t *__restrict__ row_last)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    const uint64_t key = cell_key[c];
    const uint64_t r = key / ncx;
    cell_row[c] = (uint32_t)r;
    cell_ix[c] = (int)(key - r * ncx);
    if (c == 0 || cell_key[c - 1] / ncx != r) row_first[r] = c;
    if (c == ncells - 1 || cell_key[c + 1] / ncx != r) row_last[r] = c + 1;
}

__global__ void __launch_bounds__(128, 12) k_link_thread(LinkParams P, const uint32_t *__restrict__ cell_row, const int *__restrict__ cell_ix,
                                                     const uint32_t *__restrict__ row_first, const uint32_t *__restrict__ row_last, int reach,
                                                     LinkQueue Q)
{
    const uint32_t cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= P.ncells) return;
    const uint32_t row = cell_row[cell];
    const int ix = cell_ix[cell];
    const int iz = (int)(row / (uint32_t)P.ncy);
    const int iy = (int)(row - (uint32_t)iz * (uint32_t)P.ncy);
    const uint32_t a0 = P.cell_start[cell], a1 = P.cell_start[cell + 1];
    const uint32_t na = a1 - a0;
    uint32_t ra = uf_find(P.parent, cell);
    // own row first (dx > 0), then dz = 0 / dy > 0, then dz = 1, 2
    for (int role = 0; role < 13; role++) {
        int dy, dz;
        if (role == 0) { dy = 0; dz = 0; }
        else if (role < 3) { dz = 0; dy = role; }
        else if (role < 8) { dz = 1; dy = role - 5; }
        else { dz = 2; dy = role - 10; }
        if (abs(dy) > reach || dz > reach) continue;
        int ny = iy + dy, nz = iz + dz, sy = 0, sz = 0;
        if (ny < 0) { if (!P.pery) continue; ny += P.ncy; sy = 1; }
        else if (ny >= P.ncy) { if (!P.pery) continue; ny -= P.ncy; sy = -1; }
        if (nz >= P.ncz) { if (!P.perz) continue; nz -= P.ncz; sz = -1; }
        const uint32_t r = (uint32_t)nz * (uint32_t)P.ncy + (uint32_t)ny;
        const uint32_t rf = row_first[r];
        if (rf == ROW_NONE) continue;
        const uint32_t rl = row_last[r];
        const int x0 = role == 0 ? ix + 1 : ix - reach, x1 = ix + reach; // unwrapped interval
        for (int seg = 0; seg < 2; seg++) {
            int xlo = x0, xhi = x1, xoff = 0, sx = 0;
            if (seg == 0) { xlo = max(xlo, 0); xhi = min(xhi, P.ncx - 1); }
            else {
                if (!P.perx) break;
                if (x0 < 0) { xhi = min(x1, -1) + P.ncx; xlo = x0 + P.ncx; xoff = -P.ncx; sx = 1; }
                else if (x1 >= P.ncx) { xlo = max(x0, P.ncx) - P.ncx; xhi = x1 - P.ncx; xoff = P.ncx; sx = -1; }
                else break;
            }
            if (xlo > xhi) continue;
            uint32_t c;
            if (role == 0 && seg == 0) c = cell + 1;
            else {
                uint32_t lo = rf, hi = rl;
                while (lo < hi) {
                    const uint32_t mid = (lo + hi) >> 1;
                    if (__ldg(&cell_ix[mid]) < xlo) lo = mid + 1; else hi = mid;
                }
                c = lo;
            }
            const int sc = (sx + 1) | ((sy + 1) << 2) | ((sz + 1) << 4);
            const bool wrapped = sc != (1 | (1 << 2) | (1 << 4));
            const double shx = (double)sx * P.period, shy = (double)sy * P.period, shz = (double)sz * P.period;
            for (; c < rl; c++) {
                const int bx = __ldg(&cell_ix[c]);
                if (bx > xhi) break;
                const uint32_t B = c;
                const uint32_t rb = uf_find(P.parent, B);
                if (rb == ra) continue;
                const uint32_t b0 = P.cell_start[B], b1 = P.cell_start[B + 1];
                const uint32_t nb = b1 - b0;
                // inline trial: all pairs when there are few, else the first 4 x 4
                const uint32_t ta = na * nb <= 16u ? na : min(na, 4u), tb = na * nb <= 16u ? nb : min(nb, 4u);
                bool linked = false;
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
                if (linked) {
                    uf_union(P.parent, cell, B);
                    ra = uf_find(P.parent, cell);
                } else if (na * nb > 16u) {
                    const unsigned int slot = atomicAdd(Q.count, 1u);
                    if (slot < Q.capacity) {
                        Q.a[slot] = cell;
                        Q.b[slot] = B;
                        Q.meta[slot] = (uint32_t)sc | ((uint32_t)(bx + xoff - ix + 2) << 6) | ((uint32_t)(dy + 2) << 9) | ((uint32_t)(dz + 2) << 12);
                    }
                }
            }
        }
    }
}

// one warp per undecided cell pair: the exhaustive test
__global__ void __launch_bounds__(256) k_link_queue(LinkParams P, const uint32_t *__restrict__ cell_row, const int *__restrict__ cell_ix,
                                                    LinkQueue Q, uint32_t nq)
{
    const unsigned lane = threadIdx.x & 31;
    const uint32_t warps_total = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < nq; e += warps_total) {
        const uint32_t cell = Q.a[e], B = Q.b[e], m = Q.meta[e];
        const uint32_t row = cell_row[cell];
        const int ix = cell_ix[cell];
        const int iz = (int)(row / (uint32_t)P.ncy);
        const int iy = (int)(row - (uint32_t)iz * (uint32_t)P.ncy);
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
    for (int k = 0; k < 3; k++) {
        if (g.periodic[k]) {
            g.lo[k] = 0.0;
            double m = std::ceil(g.period / cmax);
            VR_REQUIRE(m < 2.0e6, VRGPU_EUNSUPPORTED, "linking length too small for the cell key (period/b > ~1e6)");
            g.nc[k] = (int)m;
            VR_REQUIRE(g.nc[k] >= 8, VRGPU_EUNSUPPORTED, "periodic box smaller than ~5 linking lengths is not supported");
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
    const uint64_t ncell_total = (uint64_t)g.nc[0] * g.nc[1] * g.nc[2];
    const int nbits = bits_for(ncell_total - 1);
    S.grid = g;

    // ---- keys + sort ----
    DBuf<uint64_t> keys(ctx, n), keys_tmp(ctx, n);
    DBuf<uint32_t> order_tmp(ctx, n);
    S.order.alloc(ctx, n);
    DBuf<unsigned int> nbad(ctx, 1);
    VR_CUDA(cudaMemsetAsync(nbad.p, 0, sizeof(unsigned int), ctx->stream));
    {
        StageTimer t(ctx);
        { KScope k(ctx, "fof3d.cell_keys"); VR_LAUNCH(ctx, k_cell_keys, grid_for(n, T), T, 0, P.pos.p, n, g, keys.p, S.order.p, nbad.p); }
        { KScope k(ctx, "fof3d.radix_sort"); radix_sort_pairs(ctx, keys.p, S.order.p, keys_tmp.p, order_tmp.p, n, nbits); }
        S.sx.alloc(ctx, n); S.sy.alloc(ctx, n); S.sz.alloc(ctx, n);
        { KScope k(ctx, "fof3d.gather_pos"); VR_LAUNCH(ctx, k_gather_pos, grid_for(n, T), T, 0, P.pos.p, S.order.p, n, S.sx.p, S.sy.p, S.sz.p); }
        // cells
        DBuf<uint32_t> flags(ctx, n), excl(ctx, n), total(ctx, 1);
        { KScope k(ctx, "fof3d.cell_table"); VR_LAUNCH(ctx, k_flag_heads, grid_for(n, T), T, 0, keys.p, n, flags.p);
        exclusive_scan_u32(ctx, flags.p, excl.p, n, total.p); }
        uint32_t hn = 0;
        unsigned int hbad = 0;
        VR_CUDA(cudaMemcpyAsync(&hn, total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(&hbad, nbad.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        VR_REQUIRE(hbad == 0, VRGPU_EINVAL, "positions are NaN or outside [0, period]");
        S.ncells = hn;
        S.cell_of.alloc(ctx, n);
        S.cell_start.alloc(ctx, (size_t)hn + 1);
        S.cell_key.alloc(ctx, hn);
        { KScope k(ctx, "fof3d.cell_table"); VR_LAUNCH(ctx, k_fill_cells, grid_for(n, T), T, 0, keys.p, flags.p, excl.p, n, S.cell_of.p, S.cell_start.p, S.cell_key.p, hn); }
        ctx->stats.sort_ms += t.stop();
    }
    keys.reset(); keys_tmp.reset(); order_tmp.reset();
    const uint32_t ncells = S.ncells;
    ctx->stats.n_cells = ncells;

    // ---- hash + link ----
    DBuf<uint32_t> parent(ctx, ncells);
    {
        StageTimer t(ctx);
        const uint64_t nrows = (uint64_t)g.nc[1] * (uint64_t)g.nc[2];
        const bool use_rows = nrows <= (1ull << 29);
        size_t h = 1;
        if (!use_rows) while (h < 2 * (size_t)ncells) h <<= 1;
        DBuf<uint64_t> hkeys(ctx, h);
        DBuf<uint32_t> hvals(ctx, h);
        if (use_rows) {
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
        // one warp per cell, grid a multiple of the SM count
        unsigned warps_needed = ncells;
        unsigned blocks = std::min<unsigned>((warps_needed + 7) / 8, (unsigned)ctx->sm_count * 32u);
        blocks = std::max(blocks, 1u);
        if (use_rows) {
            // dense row table (L2 resident at 256^3) + thread-per-cell linking, undecided dense pairs finished by warps
            DBuf<uint32_t> row_first(ctx, nrows), row_last(ctx, nrows), cell_row(ctx, ncells);
            DBuf<int> cell_ix(ctx, ncells);
            const uint32_t qcap = (uint32_t)std::min<size_t>((size_t)8 * ncells + 1024, (size_t)0x7fffffff);
            DBuf<uint32_t> qa(ctx, qcap), qb(ctx, qcap), qm(ctx, qcap);
            DBuf<unsigned int> qcount(ctx, 1);
            LinkQueue Q{qa.p, qb.p, qm.p, qcount.p, qcap};
            KScope k(ctx, "fof3d.link_cells");
            VR_CUDA(cudaMemsetAsync(row_first.p, 0xff, sizeof(uint32_t) * nrows, ctx->stream));
            VR_LAUNCH(ctx, k_cell_coords, grid_for(ncells, T), T, 0, S.cell_key.p, ncells, (uint64_t)g.nc[0], cell_row.p, cell_ix.p, row_first.p,
                      row_last.p);
            for (int reach = 1; reach <= 2; reach++) {
                VR_CUDA(cudaMemsetAsync(qcount.p, 0, sizeof(unsigned int), ctx->stream));
                VR_LAUNCH(ctx, k_link_thread, grid_for(ncells, 128), 128, 0, L, cell_row.p, cell_ix.p, row_first.p, row_last.p, reach, Q);
                unsigned int nq = 0;
                VR_CUDA(cudaMemcpyAsync(&nq, qcount.p, sizeof(nq), cudaMemcpyDeviceToHost, ctx->stream));
                VR_CUDA(cudaStreamSynchronize(ctx->stream));
                if (nq > qcap) {
                    // queue overflow (pathological density): the warp-per-cell kernel redoes this reach exhaustively
                    VR_LAUNCH(ctx, k_link_rows, blocks, 256, 0, L, row_first.p, row_last.p, reach);
                } else if (nq) {
                    const unsigned qblocks = std::max(1u, std::min<unsigned>((nq + 7) / 8, (unsigned)ctx->sm_count * 32u));
                    VR_LAUNCH(ctx, k_link_queue, qblocks, 256, 0, L, cell_row.p, cell_ix.p, Q, nq);
                }
                ctx->stats.n_link_queue += nq;
            }
        } else {
            KScope k(ctx, "fof3d.link_cells");
            VR_LAUNCH(ctx, k_link_cells, blocks, 256, 0, L, 0, 13);
            VR_LAUNCH(ctx, k_link_cells, blocks, 256, 0, L, 13, 62);
        }
        S.root.alloc(ctx, ncells);
        { KScope k(ctx, "fof3d.flatten"); VR_LAUNCH(ctx, k_flatten, grid_for(ncells, T), T, 0, parent.p, ncells, S.root.p); }
        ctx->stats.link_ms += t.stop();
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
    if (i >= n) return;
    long long g = pfof[i];
    if (g > 0) atomicMin(&minpid[g], (unsigned long long)pid[i]);
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
