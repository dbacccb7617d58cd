// velociraptor-stf_b200/csrc/ufhash.cuh -- device helpers shared by the 3D and 6D linkers: open-addressing hash table
// (cell key -> cell id) and the lock-free union-find (atomicCAS hooking of the larger root under the smaller, path halving).
#pragma once
#include "common.cuh"

namespace vrgpu {

constexpr uint64_t HASH_EMPTY = ~0ull;

__host__ __device__ __forceinline__ uint64_t hash64(uint64_t k)
{
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdULL;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ULL;
    k ^= k >> 33;
    return k;
}

static __global__ void k_hash_fill(uint64_t *__restrict__ hkeys, size_t h)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < h) hkeys[i] = HASH_EMPTY;
}

// inserts every key and initialises parent[c] = c (when parent != nullptr)
static __global__ void k_hash_insert(const uint64_t *__restrict__ cell_key, uint32_t ncells, uint64_t *__restrict__ hkeys,
                                     uint32_t *__restrict__ hvals, uint32_t hmask, uint32_t *__restrict__ parent)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    if (parent) parent[c] = c;
    uint64_t key = cell_key[c];
    uint32_t slot = (uint32_t)hash64(key) & hmask;
    for (;;) {
        unsigned long long old = atomicCAS((unsigned long long *)&hkeys[slot], (unsigned long long)HASH_EMPTY, (unsigned long long)key);
        if (old == HASH_EMPTY) { hvals[slot] = c; return; }
        slot = (slot + 1) & hmask;
    }
}

__device__ __forceinline__ int hash_lookup(const uint64_t *__restrict__ hkeys, const uint32_t *__restrict__ hvals, uint32_t hmask, uint64_t key)
{
    uint32_t slot = (uint32_t)hash64(key) & hmask;
    for (;;) {
        uint64_t k = __ldg(&hkeys[slot]);
        if (k == key) return (int)__ldg(&hvals[slot]);
        if (k == HASH_EMPTY) return -1;
        slot = (slot + 1) & hmask;
    }
}

// roots always carry the smallest index of their set
__device__ __forceinline__ uint32_t uf_find(uint32_t *parent, uint32_t x)
{
    volatile uint32_t *vp = parent;
    uint32_t p = vp[x];
    while (p != x) {
        uint32_t gp = vp[p];
        if (gp != p) parent[x] = gp; // path halving; a stale value is still an ancestor
        x = p;
        p = gp;
    }
    return x;
}

__device__ __forceinline__ void uf_union(uint32_t *parent, uint32_t a, uint32_t b)
{
    for (;;) {
        a = uf_find(parent, a);
        b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { uint32_t t = a; a = b; b = t; }
        // a > b: hook the larger root under the smaller one
        uint32_t old = atomicCAS(&parent[a], a, b);
        if (old == a) return;
    }
}

static __global__ void k_uf_flatten(uint32_t *__restrict__ parent, uint32_t n, uint32_t *__restrict__ root)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    root[c] = uf_find(parent, c);
}

static __global__ void k_uf_init(uint32_t *__restrict__ parent, uint32_t n)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) parent[c] = c;
}

static inline int bits_for_u64v(uint64_t v)
{
    int b = 0;
    while (v) { b++; v >>= 1; }
    return b < 1 ? 1 : b;
}

// minimum of a 64-bit value over the lanes named in `peers` (every lane of the group passes the same mask, as returned
// by __match_any_sync)
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned peers, unsigned long long v)
{
    const unsigned hi = (unsigned)(v >> 32), lo = (unsigned)v;
    const unsigned mh = __reduce_min_sync(peers, hi);
    const unsigned ml = __reduce_min_sync(peers, hi == mh ? lo : 0xffffffffu);
    return ((unsigned long long)mh << 32) | ml;
}

} // namespace vrgpu
