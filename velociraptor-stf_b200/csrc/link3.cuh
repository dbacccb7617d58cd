// velociraptor-stf_b200/csrc/link3.cuh -- neighbour enumeration of the 3D cell linker (fof3d.cu), shared with its host emulation.
//
// The FOF grid has FINE cells of edge c < b/sqrt(3) (two particles of one cell are always linked), so a cell can only link to
// the cells within +-2 in every direction.  Cells are grouped into BLOCKS of 4 x 4 x 4: the sort key is
// (block index, cell inside the block), so the cells of a block are contiguous and ordered by their 6-bit local index, and
// the +-2 neighbourhood of any cell lies in exactly 2 blocks per axis -- its own and ONE other (the lower one for local
// coordinates 0, 1, the upper one for 2, 3): 8 blocks.  Per non-empty block the linker keeps a 64-bit occupancy mask and the
// index of its first cell; a bitmap over all blocks with a running count per 64-bit word (rank) finds a neighbouring block
// with two independent loads, and
//     candidates = occupancy mask of that block  &  table[local index][which of the 8 blocks]
// gives the neighbour cells of the half space (each unordered cell pair is visited once, from its lower cell) without a
// search: cell index = first cell of the block + popcount(mask below the bit).
// Everything here is plain integer code that compiles for the host too: tests/host/link3_emul.cpp runs it against a
// brute-force enumeration of the 62 half-space offsets and, as a serial linker, against the oracle.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define L3_HD __host__ __device__ __forceinline__
#else
#define L3_HD static inline
#endif

namespace vrgpu {
namespace link3 {

struct Geom {
    int ncx, ncy, ncz;    // fine cells per axis (a multiple of 4 on periodic axes)
    int nbx, nby, nbz;    // blocks per axis = ceil(nc / 4)
    int perx, pery, perz; // per axis: the grid wraps
};

// ---- sort key: (block, local) --------------------------------------------------------------------------------------
L3_HD uint64_t block_linear(const Geom &g, int bx, int by, int bz)
{
    return ((uint64_t)bz * (uint64_t)g.nby + (uint64_t)by) * (uint64_t)g.nbx + (uint64_t)bx;
}
L3_HD uint64_t cell_key(const Geom &g, int ix, int iy, int iz)
{
    const unsigned loc = (unsigned)(((iz & 3) << 4) | ((iy & 3) << 2) | (ix & 3));
    return (block_linear(g, ix >> 2, iy >> 2, iz >> 2) << 6) | loc;
}
L3_HD void key_cell(const Geom &g, uint64_t key, int &ix, int &iy, int &iz)
{
    const unsigned loc = (unsigned)(key & 63u);
    uint64_t b = key >> 6;
    const uint64_t nxy = (uint64_t)g.nbx * (uint64_t)g.nby;
    const int bz = (int)(b / nxy);
    b -= (uint64_t)bz * nxy;
    const int by = (int)(b / (uint64_t)g.nbx);
    const int bx = (int)(b - (uint64_t)by * (uint64_t)g.nbx);
    ix = (bx << 2) | (int)(loc & 3u);
    iy = (by << 2) | (int)((loc >> 2) & 3u);
    iz = (bz << 2) | (int)(loc >> 4);
}
// block coordinates packed into 64 bits (21 bits each)
L3_HD uint64_t pack_xyz(int bx, int by, int bz) { return (uint64_t)bx | ((uint64_t)by << 21) | ((uint64_t)bz << 42); }
L3_HD void unpack_xyz(uint64_t v, int &bx, int &by, int &bz)
{
    bx = (int)(v & 0x1fffffu); by = (int)((v >> 21) & 0x1fffffu); bz = (int)((v >> 42) & 0x1fffffu);
}

// ---- per-axis 4-bit masks ------------------------------------------------------------------------------------------
// cells of a block within `reach` of local coordinate l; which = 0: the own block, 1: the other block (below for l < 2,
// above for l >= 2)
L3_HD unsigned axis_reach(int l, int which, int reach)
{
    int lo = l - reach, hi = l + reach;
    if (which == 0) { lo = lo < 0 ? 0 : lo; hi = hi > 3 ? 3 : hi; }
    else if (l < 2) { if (lo > -1) return 0u; lo += 4; hi = 3; }
    else { if (hi < 4) return 0u; lo = 0; hi -= 4; }
    return ((1u << (hi + 1)) - 1u) & ~((1u << lo) - 1u);
}
// ... of those, the cells with a strictly larger global coordinate / with the same coordinate
L3_HD unsigned axis_greater(int l, int which, int reach)
{
    const unsigned r = axis_reach(l, which, reach);
    if (which == 0) return r & ~((2u << l) - 1u);
    return l < 2 ? 0u : r;
}
L3_HD unsigned axis_equal(int l, int which) { return which == 0 ? (1u << l) : 0u; }

L3_HD uint64_t cube(unsigned mx, unsigned my, unsigned mz)
{
    const unsigned sy = (my & 1u) | ((my & 2u) << 3) | ((my & 4u) << 6) | ((my & 8u) << 9);
    const uint64_t sz = (uint64_t)(mz & 1u) | ((uint64_t)(mz & 2u) << 15) | ((uint64_t)(mz & 4u) << 30) | ((uint64_t)(mz & 8u) << 45);
    return (uint64_t)(mx * sy) * sz;
}

// cells (bit = x + 4 y + 16 z inside the block) of block j (bit 0 / 1 / 2 of j: the other block along x / y / z) that lie within
// `reach` of the cell with local index loc AND in its forward half space: dz > 0, or dz == 0 and dy > 0, or dz == dy == 0
// and dx > 0 in unwrapped grid coordinates
L3_HD uint64_t forward_mask(unsigned loc, unsigned j, int reach)
{
    const int lx = (int)(loc & 3u), ly = (int)((loc >> 2) & 3u), lz = (int)(loc >> 4);
    const int jx = (int)(j & 1u), jy = (int)((j >> 1) & 1u), jz = (int)(j >> 2);
    const unsigned rx = axis_reach(lx, jx, reach), ry = axis_reach(ly, jy, reach);
    const unsigned gx = axis_greater(lx, jx, reach), gy = axis_greater(ly, jy, reach), gz = axis_greater(lz, jz, reach);
    const unsigned ey = axis_equal(ly, jy), ez = axis_equal(lz, jz);
    return cube(rx, ry, gz) | cube(rx, gy, ez) | cube(gx, ey, ez);
}

// table[pass][loc][j]: pass 0 = the 13 adjacent half-space cells (reach 1), pass 1 = the 49 of the second shell
constexpr int TABLE_WORDS = 2 * 64 * 8;
inline void build_table(uint64_t *t)
{
    for (unsigned loc = 0; loc < 64; loc++)
        for (unsigned j = 0; j < 8; j++) {
            const uint64_t r1 = forward_mask(loc, j, 1), r2 = forward_mask(loc, j, 2);
            t[(0 * 64 + loc) * 8 + j] = r1;
            t[(1 * 64 + loc) * 8 + j] = r2 & ~r1;
        }
}

// ---- the other block along one axis ----------------------------------------------------------------------------------
// b: block coordinate, l: local coordinate, which: 0 own / 1 other.  Returns false when the block lies outside an open
// grid.  shift = +1: the neighbour was reached through the lower face (its coordinates are one period ABOVE what the
// pair predicate needs, so the target is shifted by +period), -1: through the upper face -- the convention of the image
// code `sc` in fof3d.cu.  doff = offset of the neighbour block in unwrapped block coordinates (-1, 0, +1).
L3_HD bool axis_block(int b, int l, int which, int nb, int periodic, int &nbk, int &shift, int &doff)
{
    doff = which ? (l < 2 ? -1 : 1) : 0;
    nbk = b + doff;
    shift = 0;
    if (nbk < 0) { if (!periodic) return false; nbk += nb; shift = 1; }
    else if (nbk >= nb) { if (!periodic) return false; nbk -= nb; shift = -1; }
    return true;
}

// ---- the 8 blocks around a cell ----------------------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
#define L3_LD64(p) __ldg(p)
#define L3_LD32(p) __ldg(p)
#define L3_POPC64(v) __popcll(v)
#else
#define L3_LD64(p) (*(p))
#define L3_LD32(p) (*(p))
#define L3_POPC64(v) __builtin_popcountll(v)
#endif

struct Tables {
    const unsigned long long *bitmap; // one bit per block of the grid, block_linear order
    const uint32_t *prefix;           // per 64-bit word: number of non-empty blocks before it
    const unsigned long long *blk_occ; // per non-empty block (by rank): occupancy of its 64 cells
    const uint32_t *blk_cell0;        // per non-empty block: index of its first cell
};

// one 64-bit record per cell: block coordinates (19 bits each) and the local index
L3_HD uint64_t pack_cell(int bx, int by, int bz, unsigned loc)
{
    return (uint64_t)bx | ((uint64_t)by << 19) | ((uint64_t)bz << 38) | ((uint64_t)loc << 57);
}
L3_HD void unpack_cell(uint64_t v, int &bx, int &by, int &bz, unsigned &loc)
{
    bx = (int)(v & 0x7ffffu); by = (int)((v >> 19) & 0x7ffffu); bz = (int)((v >> 38) & 0x7ffffu); loc = (unsigned)(v >> 57) & 63u;
}

// For the cell with local index loc in the block at (bx, by, bz): rank and occupancy of the 8 blocks its neighbourhood touches
// (occ[j] == 0: empty, outside an open grid, or wholly behind the cell) and the image code of each (6 bits per j in `codes`:
// (sx+1) | (sy+1) << 2 | (sz+1) << 4).  Two levels of independent loads: 8 bitmap words, then rank + occupancy of the blocks
// that exist (j = 0 is the cell's own block).
L3_HD void lookup_blocks(const Geom &g, const Tables &T, int bx, int by, int bz, unsigned loc, uint32_t rank[8], unsigned long long occ[8],
                         unsigned long long &codes)
{
    const int lx = (int)(loc & 3u), ly = (int)((loc >> 2) & 3u), lz = (int)(loc >> 4);
    int n1, s1[3], d;
    bool ok1[3];
    // linear index = ax[jx] + ay[jy] + az[jz]
    uint64_t ax[2], ay[2], az[2];
    const uint64_t nxy = (uint64_t)g.nbx * (uint64_t)g.nby;
    ax[0] = (uint64_t)bx; ay[0] = (uint64_t)by * (uint64_t)g.nbx; az[0] = (uint64_t)bz * nxy;
    ok1[0] = axis_block(bx, lx, 1, g.nbx, g.perx, n1, s1[0], d); ax[1] = (uint64_t)n1;
    ok1[1] = axis_block(by, ly, 1, g.nby, g.pery, n1, s1[1], d); ay[1] = (uint64_t)n1 * (uint64_t)g.nbx;
    ok1[2] = axis_block(bz, lz, 1, g.nbz, g.perz, n1, s1[2], d); az[1] = (uint64_t)n1 * nxy;
    // the other block along z lies behind the cell when it is the lower one: nothing of it is in the forward half space
    if (lz < 2) ok1[2] = false;
    unsigned long long word[8];
    uint64_t blin[8];
    codes = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int j = 0; j < 8; j++) {
        const int jx = j & 1, jy = (j >> 1) & 1, jz = j >> 2;
        const bool ok = (!jx || ok1[0]) && (!jy || ok1[1]) && (!jz || ok1[2]);
        blin[j] = ax[jx] + ay[jy] + az[jz];
        word[j] = ok ? L3_LD64(T.bitmap + (blin[j] >> 6)) : 0ull;
        codes |= (unsigned long long)(((jx ? s1[0] : 0) + 1) | (((jy ? s1[1] : 0) + 1) << 2) | (((jz ? s1[2] : 0) + 1) << 4)) << (6 * j);
    }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int j = 0; j < 8; j++) {
        const unsigned bit = (unsigned)(blin[j] & 63u);
        const bool there = ((word[j] >> bit) & 1ull) != 0; // word == 0 for the blocks that were not looked at
        rank[j] = there ? L3_LD32(T.prefix + (blin[j] >> 6)) + (uint32_t)L3_POPC64(word[j] & ((1ull << bit) - 1ull)) : 0u;
        occ[j] = there ? L3_LD64(T.blk_occ + rank[j]) : 0ull;
    }
}

// the cell behind bit `bit` of block `rank` (occupancy occ)
L3_HD uint32_t cell_of_bit(const Tables &T, uint32_t rank, unsigned long long occ, unsigned bit)
{
    return L3_LD32(T.blk_cell0 + rank) + (uint32_t)L3_POPC64(occ & ((1ull << bit) - 1ull));
}

// offset (in fine cells, unwrapped) from the cell with local index loc to the cell behind bit `bit` of its block j
L3_HD void cell_offset(unsigned loc, unsigned j, unsigned bit, int &ddx, int &ddy, int &ddz)
{
    const int lx = (int)(loc & 3u), ly = (int)((loc >> 2) & 3u), lz = (int)(loc >> 4);
    const int ox = (j & 1u) ? (lx < 2 ? -4 : 4) : 0, oy = (j & 2u) ? (ly < 2 ? -4 : 4) : 0, oz = (j & 4u) ? (lz < 2 ? -4 : 4) : 0;
    ddx = ox + (int)(bit & 3u) - lx;
    ddy = oy + (int)((bit >> 2) & 3u) - ly;
    ddz = oz + (int)(bit >> 4) - lz;
}

} // namespace link3
} // namespace vrgpu
