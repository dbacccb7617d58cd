// velociraptor-stf_b200/csrc/multigpu.cu -- slab-decomposed multi-GPU structure finding (one process per GPU, NCCL).
//
// Replaces the reference's MPI path through SearchFullSet (src/search.cxx:242-347): MPIInitialDomainDecomposition /
// MPIInDomain (mpiroutines.cxx:49,259), MPIGetExportNum + MPIBuildParticleExportList (:783,:867: ship every particle
// within one linking length of a neighbouring domain), the iterated MPILinkAcross / MPIUpdateExportList fix point
// (:2387,:2318, search.cxx:301-313), MPIGroupExchange (:2548: every group is shipped whole to ONE rank before 6DFOF /
// unbinding) and MPICompileGroups (:2818).
//
// Device design:
//   * rank r owns the x-slab [slab_lo, slab_hi) of the periodic box; particles within b of a face are copied to the
//     neighbouring rank as GHOSTS (ncclSend/ncclRecv; the periodic wrap is applied to the ghost's x so the pair
//     predicate sees exactly the operands the single-GPU path forms for a wrapped pair);
//   * every rank links owned + ghost particles with the ordinary cell-grid linker (x open, y/z periodic);
//   * each ghost yields one EDGE (my component, the owner's component of the same particle).  All edges are
//     all-gathered (they are O(N^(2/3))) and every rank runs the same small device union-find over the boundary
//     components -- one round instead of the reference's iterate-until-no-change loop;
//   * group sizes / min pid: local for interior components, ncclAllReduce(SUM/MIN) over the replicated boundary table;
//     all candidate groups are all-gathered and ordered by (size desc, min pid asc): ids are GLOBAL and identical to the
//     single-GPU ids;
//   * for 6DFOF / unbinding every boundary group is consolidated on the rank that holds its root label (members sent
//     with grouped ncclSend/ncclRecv), the ordinary single-GPU stages run on owned + imported particles, the surviving
//     groups are renumbered globally and the labels of imported particles travel back to their owners.
#include "groups.cuh"
#include "stages.cuh"
#include "ufhash.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <memory>
#include <cmath>
#include <numeric>

namespace vrgpu {

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi &nccl_api()
{
    static NcclApi api;
    if (api.lib) return api;
    // inside a torch process this resolves to the already loaded (bundled) libnccl.so.2; otherwise to the system one
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.lib) break;
    }
    VR_REQUIRE(api.lib != nullptr, VRGPU_EUNSUPPORTED, "libnccl.so.2 not found: multi-GPU runs need NCCL");
#define VR_NCCL_SYM(field, sym)                                                       \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, sym));           \
    VR_REQUIRE(api.field != nullptr, VRGPU_EUNSUPPORTED, "NCCL symbol missing: " sym)
    VR_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
    VR_NCCL_SYM(CommInitRank, "ncclCommInitRank");
    VR_NCCL_SYM(CommDestroy, "ncclCommDestroy");
    VR_NCCL_SYM(AllGather, "ncclAllGather");
    VR_NCCL_SYM(AllReduce, "ncclAllReduce");
    VR_NCCL_SYM(Send, "ncclSend");
    VR_NCCL_SYM(Recv, "ncclRecv");
    VR_NCCL_SYM(GroupStart, "ncclGroupStart");
    VR_NCCL_SYM(GroupEnd, "ncclGroupEnd");
    VR_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef VR_NCCL_SYM
    return api;
}

#define VR_NCCL(call)                                                                                         \
    do {                                                                                                      \
        ncclResult_t r__ = (call);                                                                            \
        if (r__ != ncclSuccess)                                                                               \
            throw ::vrgpu::Error(VRGPU_ECUDA, std::string(#call) + ": " + nccl_api().GetErrorString(r__));    \
    } while (0)

struct MultiGpu {
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0;
};

void mg_destroy(vrgpu_ctx *ctx)
{
    if (!ctx || !ctx->mg) return;
    if (ctx->mg->comm) nccl_api().CommDestroy(ctx->mg->comm);
    delete ctx->mg;
    ctx->mg = nullptr;
}

void mg_unique_id(char *out128)
{
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    VR_NCCL(nccl_api().GetUniqueId(&id));
    std::memcpy(out128, &id, 128);
}

void mg_init(vrgpu_ctx *ctx, int nranks, int rank, const char *id128)
{
    VR_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, VRGPU_EINVAL, "bad rank / nranks");
    mg_destroy(ctx);
    ctx->mg = new MultiGpu;
    ctx->mg->nranks = nranks;
    ctx->mg->rank = rank;
    if (nranks > 1) {
        ncclUniqueId id;
        std::memcpy(&id, id128, 128);
        VR_NCCL(nccl_api().CommInitRank(&ctx->mg->comm, nranks, id, rank));
    }
}

int mg_nranks(const vrgpu_ctx *ctx) { return ctx->mg ? ctx->mg->nranks : 1; }
int mg_rank(const vrgpu_ctx *ctx) { return ctx->mg ? ctx->mg->rank : 0; }

namespace {

constexpr uint32_t MG_NONE = 0xffffffffu;

// ---- small collective helpers (device buffers, context stream) -------------------------------------------
// all-gather one fixed-size record per rank and bring it to the host
template <typename T> std::vector<T> allgather_host(vrgpu_ctx *ctx, const T *host_rec, int per_rank)
{
    MultiGpu &M = *ctx->mg;
    DBuf<T> in(ctx, (size_t)per_rank), out(ctx, (size_t)per_rank * M.nranks);
    VR_CUDA(cudaMemcpyAsync(in.p, host_rec, sizeof(T) * per_rank, cudaMemcpyHostToDevice, ctx->stream));
    VR_NCCL(nccl_api().AllGather(in.p, out.p, sizeof(T) * per_rank, ncclUint8, M.comm, ctx->stream));
    std::vector<T> h((size_t)per_rank * M.nranks);
    VR_CUDA(cudaMemcpyAsync(h.data(), out.p, sizeof(T) * h.size(), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return h;
}

// variable-length all-gather of 8-byte words: returns the concatenation (rank order) and the per-rank counts
void allgatherv_u64(vrgpu_ctx *ctx, const uint64_t *mine, size_t count, DBuf<uint64_t> &all, std::vector<size_t> &counts)
{
    MultiGpu &M = *ctx->mg;
    unsigned long long c = count;
    std::vector<unsigned long long> cs = allgather_host<unsigned long long>(ctx, &c, 1);
    counts.assign(cs.begin(), cs.end());
    size_t mx = 0, tot = 0;
    for (size_t v : counts) { mx = std::max(mx, v); tot += v; }
    all.alloc(ctx, tot ? tot : 1);
    if (mx == 0) return;
    DBuf<uint64_t> in(ctx, mx), out(ctx, mx * M.nranks);
    if (count) VR_CUDA(cudaMemcpyAsync(in.p, mine, count * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    VR_NCCL(nccl_api().AllGather(in.p, out.p, mx * 8, ncclUint8, M.comm, ctx->stream));
    size_t o = 0;
    for (int r = 0; r < M.nranks; r++) {
        if (counts[r]) VR_CUDA(cudaMemcpyAsync(all.p + o, out.p + (size_t)r * mx, counts[r] * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        o += counts[r];
    }
    VR_CUDA(cudaStreamSynchronize(ctx->stream)); // `in`/`out` go back to the pool
}

} // namespace
} // namespace vrgpu

namespace vrgpu {
namespace {

// ---- ghosts ------------------------------------------------------------------------------------------------
__global__ void k_mg_flag_ghost(const double *__restrict__ pos, size_t n, double xl, double xr, uint32_t *__restrict__ fl, uint32_t *__restrict__ fr)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double x = pos[3 * i];
    fl[i] = x < xl ? 1u : 0u;
    fr[i] = x >= xr ? 1u : 0u;
}
__global__ void k_mg_pack_ghost(const double *__restrict__ pos, const long long *__restrict__ pid, const uint32_t *__restrict__ flag,
                                const uint32_t *__restrict__ excl, size_t n, double shift, uint32_t *__restrict__ idx, double *__restrict__ opos,
                                long long *__restrict__ opid)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !flag[i]) return;
    const uint32_t k = excl[i];
    idx[k] = (uint32_t)i;
    opos[3 * (size_t)k] = shift != 0.0 ? __dadd_rn(pos[3 * i], shift) : pos[3 * i];
    opos[3 * (size_t)k + 1] = pos[3 * i + 1];
    opos[3 * (size_t)k + 2] = pos[3 * i + 2];
    opid[k] = pid ? pid[i] : (long long)i;
}
// component (root cell) of every particle of the extended set, by original index
__global__ void k_mg_comp(const uint32_t *__restrict__ order, const uint32_t *__restrict__ cell_of, const uint32_t *__restrict__ root, size_t n,
                          uint32_t *__restrict__ comp)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) comp[order[k]] = root[cell_of[k]];
}
__global__ void k_mg_send_labels(const uint32_t *__restrict__ idx, const uint32_t *__restrict__ comp, uint32_t cnt, uint64_t rank, uint64_t *__restrict__ out)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < cnt) out[k] = (rank << 32) | comp[idx[k]];
}
// edges[2j] = my component of ghost j, edges[2j+1] = the owner's component of the same particle
__global__ void k_mg_edges(const uint32_t *__restrict__ comp, size_t n_own, const uint64_t *__restrict__ recv, uint32_t cnt, uint64_t rank,
                           uint64_t *__restrict__ edges)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= cnt) return;
    edges[2 * (size_t)j] = (rank << 32) | comp[n_own + j];
    edges[2 * (size_t)j + 1] = recv[j];
}

// ---- replicated union-find over the boundary components -------------------------------------------------------
__global__ void k_mg_iota(uint32_t *a, uint32_t n) { uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) a[i] = i; }
__global__ void k_mg_heads(const uint64_t *__restrict__ keys, uint32_t n, uint32_t *__restrict__ flags)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) flags[k] = (k == 0 || keys[k] != keys[k - 1]) ? 1u : 0u;
}
// cidx[original endpoint position] = compact id; U[compact id] = label
__global__ void k_mg_compact_ids(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ vals, const uint32_t *__restrict__ flags,
                                 const uint32_t *__restrict__ excl, uint32_t n, uint32_t *__restrict__ cidx, uint64_t *__restrict__ U)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t c = excl[k] - (flags[k] ? 0u : 1u);
    cidx[vals[k]] = c;
    if (flags[k]) U[c] = keys[k];
}
__global__ void k_mg_union(const uint32_t *__restrict__ cidx, uint32_t nedges, uint32_t *__restrict__ parent)
{
    uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < nedges) uf_union(parent, cidx[2 * e], cidx[2 * e + 1]);
}

// ---- sizes ---------------------------------------------------------------------------------------------------
__global__ void k_mg_comp_stats(const uint32_t *__restrict__ comp, const long long *__restrict__ pid, size_t n_own, uint32_t *__restrict__ csize,
                                unsigned long long *__restrict__ cminpid)
{
    // neighbours in the caller's array are usually neighbours in space (a snapshot is written domain by domain): one atomic per
    // (warp, component) instead of one per particle -- a 10^6-member component was 10^6 atomics on one address
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n_own;
    const uint32_t c = valid ? comp[i] : 0xffffffffu;
    const unsigned long long v = valid ? (pid ? (unsigned long long)pid[i] : (unsigned long long)i) : ~0ull;
    const unsigned peers = __match_any_sync(0xffffffffu, c);
    const unsigned long long mn = warp_min_u64(peers, v);
    if (valid && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) {
        atomicAdd(&csize[c], (uint32_t)__popc(peers));
        atomicMin(&cminpid[c], mn);
    }
}
// per boundary entry owned by this rank: its local size / min pid; bidx[cell] = compact id
__global__ void k_mg_boundary_contrib(const uint64_t *__restrict__ U, uint32_t nU, uint64_t rank, const uint32_t *__restrict__ csize,
                                      const unsigned long long *__restrict__ cminpid, unsigned long long *__restrict__ usize,
                                      unsigned long long *__restrict__ uminpid, uint32_t *__restrict__ bidx)
{
    uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= nU) return;
    const uint64_t lab = U[u];
    if ((lab >> 32) == rank) {
        const uint32_t c = (uint32_t)lab;
        usize[u] = csize[c];
        uminpid[u] = cminpid[c];
        bidx[c] = u;
    } else {
        usize[u] = 0;
        uminpid[u] = ~0ull;
    }
}
__global__ void k_mg_boundary_totals(const uint32_t *__restrict__ R, uint32_t nU, const unsigned long long *__restrict__ usize,
                                     const unsigned long long *__restrict__ uminpid, unsigned long long *__restrict__ gsz,
                                     unsigned long long *__restrict__ gmp)
{
    uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= nU) return;
    const uint32_t r = R[u];
    if (usize[u]) atomicAdd(&gsz[r], usize[u]);
    atomicMin(&gmp[r], uminpid[u]);
}

// ---- candidate records: 3 words (size, min pid, kind<<63 | ref) ----------------------------------------------------
__global__ void k_mg_flag_local(const uint32_t *__restrict__ root, const uint32_t *__restrict__ bidx, const uint32_t *__restrict__ csize,
                                uint32_t ncells, uint32_t minsize, uint32_t *__restrict__ flags)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < ncells) flags[c] = (root[c] == c && bidx[c] == MG_NONE && csize[c] >= minsize) ? 1u : 0u;
}
__global__ void k_mg_flag_boundary(const uint32_t *__restrict__ R, const uint64_t *__restrict__ U, uint32_t nU, uint64_t rank,
                                   const unsigned long long *__restrict__ gsz, uint32_t minsize, uint32_t *__restrict__ flags)
{
    uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u < nU) flags[u] = (R[u] == u && (U[u] >> 32) == rank && gsz[u] >= minsize) ? 1u : 0u;
}
__global__ void k_mg_records_local(const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl, uint32_t ncells,
                                   const uint32_t *__restrict__ csize, const unsigned long long *__restrict__ cminpid, uint64_t *__restrict__ rec)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells || !flags[c]) return;
    uint64_t *r = rec + 3 * (size_t)excl[c];
    r[0] = csize[c]; r[1] = cminpid[c]; r[2] = c;
}
__global__ void k_mg_records_boundary(const uint32_t *__restrict__ flags, const uint32_t *__restrict__ excl, uint32_t nU,
                                      const unsigned long long *__restrict__ gsz, const unsigned long long *__restrict__ gmp,
                                      uint64_t *__restrict__ rec)
{
    uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= nU || !flags[u]) return;
    uint64_t *r = rec + 3 * (size_t)excl[u];
    r[0] = gsz[u]; r[1] = gmp[u]; r[2] = (1ull << 63) | u;
}
__global__ void k_mg_split_records(const uint64_t *__restrict__ rec, uint32_t n, uint32_t *__restrict__ size, unsigned long long *__restrict__ minpid)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    size[j] = (uint32_t)rec[3 * (size_t)j];
    minpid[j] = rec[3 * (size_t)j + 1];
}
// scatter the global ids: boundary records (any rank) -> gid_u[ref]; local records of THIS rank -> gid_cell[ref]
__global__ void k_mg_scatter_ids(const uint64_t *__restrict__ rec, const uint32_t *__restrict__ gid_of, uint32_t n, uint32_t my0, uint32_t my1,
                                 uint32_t *__restrict__ gid_u, uint32_t *__restrict__ gid_cell)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint64_t ref = rec[3 * (size_t)j + 2];
    if (ref >> 63) gid_u[(uint32_t)ref] = gid_of[j];
    else if (j >= my0 && j < my1) gid_cell[(uint32_t)ref] = gid_of[j];
}
__global__ void k_mg_write_pfof(const uint32_t *__restrict__ comp, const uint32_t *__restrict__ bidx, const uint32_t *__restrict__ R,
                                const uint32_t *__restrict__ gid_u, const uint32_t *__restrict__ gid_cell, size_t n_own, long long *__restrict__ pfof)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_own) return;
    const uint32_t c = comp[i];
    const uint32_t u = bidx[c];
    pfof[i] = u != MG_NONE ? gid_u[R[u]] : gid_cell[c];
}
__global__ void k_mg_dest_of(const long long *__restrict__ pfof, const int *__restrict__ dest_of_gid, size_t n_own, int *__restrict__ dest)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_own) dest[i] = dest_of_gid[pfof[i]];
}

} // namespace
} // namespace vrgpu

namespace vrgpu {
namespace {

struct GlobalOrder {
    DBuf<uint64_t> rec;      // gathered records (3 words each), rank order
    DBuf<uint32_t> gid_of;   // global id of every gathered record
    uint32_t total = 0, my0 = 0, my1 = 0;
    std::vector<uint32_t> rank_first; // first record of every rank (nranks + 1 entries)
};

// all-gather this rank's candidate records and order ALL of them by (size desc, min pid asc)
void mg_global_order(vrgpu_ctx *ctx, const uint64_t *myrec, size_t nmine, GlobalOrder &O)
{
    MultiGpu &M = *ctx->mg;
    const int T = 256;
    std::vector<size_t> counts;
    allgatherv_u64(ctx, myrec, 3 * nmine, O.rec, counts);
    size_t tot = 0, before = 0;
    O.rank_first.assign(M.nranks + 1, 0);
    for (int r = 0; r < M.nranks; r++) {
        if (r < M.rank) before += counts[r] / 3;
        tot += counts[r] / 3;
        O.rank_first[r + 1] = (uint32_t)std::min<size_t>(tot, 0xffffffffull);
    }
    VR_REQUIRE(tot < 0xfffffff0ull, VRGPU_EUNSUPPORTED, "more than 2^32 groups");
    O.total = (uint32_t)tot;
    O.my0 = (uint32_t)before;
    O.my1 = (uint32_t)(before + nmine);
    O.gid_of.alloc(ctx, tot ? tot : 1);
    if (tot == 0) return;
    DBuf<uint32_t> iota(ctx, tot), size(ctx, tot);
    DBuf<unsigned long long> minpid(ctx, tot);
    VR_LAUNCH(ctx, k_mg_iota, grid_for(tot, T), T, 0, iota.p, (uint32_t)tot);
    VR_LAUNCH(ctx, k_mg_split_records, grid_for(tot, T), T, 0, O.rec.p, (uint32_t)tot, size.p, minpid.p);
    order_groups(ctx, iota.p, size.p, minpid.p, (uint32_t)tot, 1u, O.gid_of.p);
}

struct StageA {
    DBuf<long long> pfof;    // global 3D group id of every owned particle
    DBuf<int> dest;          // rank that consolidates the particle's group (-1: ungrouped); filled by mg_place_groups
    long long ngroups = 0;
    // replicated on every rank, indexed by global id - 1: member count, and the rank holding ALL members (-1 when the group
    // straddles a slab face)
    std::vector<uint32_t> gsize;
    std::vector<int> ghome;
    size_t n_ghost = 0, n_edges = 0, n_boundary = 0;
};

// 3DFOF over all ranks with global ids
void mg_fof3d(vrgpu_ctx *ctx, const DeviceParticles &P, double slab_lo, double slab_hi, double b, double period, long long minsize, StageA &A)
{
    MultiGpu &M = *ctx->mg;
    const int T = 256;
    const int N = M.nranks, me = M.rank;
    const size_t n_own = P.n;
    VR_REQUIRE(period > 0, VRGPU_EUNSUPPORTED, "multi-GPU runs need a periodic box");
    VR_REQUIRE(slab_hi - slab_lo > 4.0 * b, VRGPU_EUNSUPPORTED, "slab thinner than 4 linking lengths");
    const int left = (me + N - 1) % N, right = (me + 1) % N;
    const double bm = b * (1.0 + 1e-9);

    // ---- ghosts: particles within b of a face go to the neighbour, x wrapped where the face is the box boundary ----
    KScope kall(ctx, "mg.fof3d_total");
    DBuf<uint32_t> idxL, idxR;
    DBuf<double> sposL, sposR;
    DBuf<long long> spidL, spidR;
    uint32_t nSL = 0, nSR = 0;
    {
        KScope k(ctx, "mg.ghost_pack");
        DBuf<uint32_t> fl(ctx, n_own + 1), fr(ctx, n_own + 1), el(ctx, n_own + 1), er(ctx, n_own + 1), tot(ctx, 2);
        if (n_own) VR_LAUNCH(ctx, k_mg_flag_ghost, grid_for(n_own, T), T, 0, P.pos.p, n_own, slab_lo + bm, slab_hi - bm, fl.p, fr.p);
        exclusive_scan_u32(ctx, fl.p, el.p, n_own, tot.p);
        exclusive_scan_u32(ctx, fr.p, er.p, n_own, tot.p + 1);
        uint32_t h[2];
        VR_CUDA(cudaMemcpyAsync(h, tot.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        nSL = h[0]; nSR = h[1];
        idxL.alloc(ctx, nSL + 1); sposL.alloc(ctx, 3 * (size_t)nSL + 1); spidL.alloc(ctx, nSL + 1);
        idxR.alloc(ctx, nSR + 1); sposR.alloc(ctx, 3 * (size_t)nSR + 1); spidR.alloc(ctx, nSR + 1);
        const long long *pid = P.has_pid ? P.pid.p : nullptr;
        if (n_own) {
            // to the left neighbour: rank 0's left neighbour sits at the top of the box -> +period; mirrored for the right
            VR_LAUNCH(ctx, k_mg_pack_ghost, grid_for(n_own, T), T, 0, P.pos.p, pid, fl.p, el.p, n_own, me == 0 ? period : 0.0, idxL.p, sposL.p, spidL.p);
            VR_LAUNCH(ctx, k_mg_pack_ghost, grid_for(n_own, T), T, 0, P.pos.p, pid, fr.p, er.p, n_own, me == N - 1 ? -period : 0.0, idxR.p, sposR.p, spidR.p);
        }
    }
    unsigned long long mycnt[2] = {nSL, nSR};
    std::vector<unsigned long long> allc = allgather_host<unsigned long long>(ctx, mycnt, 2);
    const uint32_t nRL = (uint32_t)allc[2 * left + 1];  // what my left neighbour sends to its right
    const uint32_t nRR = (uint32_t)allc[2 * right + 0]; // what my right neighbour sends to its left
    const size_t n_ext = n_own + nRR + nRL;
    A.n_ghost = nRR + nRL;

    // extended particle set: owned, then ghosts from the right, then ghosts from the left
    DeviceParticles X;
    X.n = n_ext;
    X.pos.alloc(ctx, 3 * n_ext + 1);
    X.pid.alloc(ctx, n_ext + 1);
    X.has_pid = true;
    {
        KScope k(ctx, "mg.ghost_exchange");
        if (n_own) VR_CUDA(cudaMemcpyAsync(X.pos.p, P.pos.p, 3 * n_own * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        if (n_own) {
            if (P.has_pid) VR_CUDA(cudaMemcpyAsync(X.pid.p, P.pid.p, n_own * sizeof(long long), cudaMemcpyDeviceToDevice, ctx->stream));
            else VR_CUDA(cudaMemsetAsync(X.pid.p, 0, n_own * sizeof(long long), ctx->stream));
        }
        VR_NCCL(nccl_api().GroupStart());
        VR_NCCL(nccl_api().Send(sposL.p, 3 * (size_t)nSL * 8, ncclUint8, left, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Send(sposR.p, 3 * (size_t)nSR * 8, ncclUint8, right, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Recv(X.pos.p + 3 * n_own, 3 * (size_t)nRR * 8, ncclUint8, right, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Recv(X.pos.p + 3 * (n_own + nRR), 3 * (size_t)nRL * 8, ncclUint8, left, M.comm, ctx->stream));
        VR_NCCL(nccl_api().GroupEnd());
        VR_NCCL(nccl_api().GroupStart());
        VR_NCCL(nccl_api().Send(spidL.p, (size_t)nSL * 8, ncclUint8, left, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Send(spidR.p, (size_t)nSR * 8, ncclUint8, right, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Recv(X.pid.p + n_own, (size_t)nRR * 8, ncclUint8, right, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Recv(X.pid.p + n_own + nRR, (size_t)nRL * 8, ncclUint8, left, M.comm, ctx->stream));
        VR_NCCL(nccl_api().GroupEnd());
    }

    // ---- local linking of owned + ghost particles (x open: the ghosts carry the wrap) ----
    Fof3dState S;
    const int axes[3] = {0, 1, 1};
    fof3d_link(ctx, X, b, period, axes, S);
    const uint32_t ncells = S.ncells;
    DBuf<uint32_t> comp(ctx, n_ext + 1);
    if (n_ext) VR_LAUNCH(ctx, k_mg_comp, grid_for(n_ext, T), T, 0, S.order.p, S.cell_of.p, S.root.p, n_ext, comp.p);

    // ---- edges: one per ghost ----
    const uint32_t nE = nRR + nRL;
    DBuf<uint64_t> edges(ctx, 2 * (size_t)nE + 1);
    {
        KScope k(ctx, "mg.label_exchange");
        DBuf<uint64_t> slabL(ctx, nSL + 1), slabR(ctx, nSR + 1), rlab(ctx, nE + 1);
        if (nSL) VR_LAUNCH(ctx, k_mg_send_labels, grid_for(nSL, T), T, 0, idxL.p, comp.p, nSL, (uint64_t)me, slabL.p);
        if (nSR) VR_LAUNCH(ctx, k_mg_send_labels, grid_for(nSR, T), T, 0, idxR.p, comp.p, nSR, (uint64_t)me, slabR.p);
        VR_NCCL(nccl_api().GroupStart());
        VR_NCCL(nccl_api().Send(slabL.p, (size_t)nSL * 8, ncclUint8, left, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Send(slabR.p, (size_t)nSR * 8, ncclUint8, right, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Recv(rlab.p, (size_t)nRR * 8, ncclUint8, right, M.comm, ctx->stream));
        VR_NCCL(nccl_api().Recv(rlab.p + nRR, (size_t)nRL * 8, ncclUint8, left, M.comm, ctx->stream));
        VR_NCCL(nccl_api().GroupEnd());
        if (nE) VR_LAUNCH(ctx, k_mg_edges, grid_for(nE, T), T, 0, comp.p, n_own, rlab.p, nE, (uint64_t)me, edges.p);
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
    }

    // ---- replicated union-find over the boundary components ----
    DBuf<uint64_t> alle, U;
    DBuf<uint32_t> R;
    uint32_t nU = 0;
    {
        KScope k(ctx, "mg.boundary_merge");
        std::vector<size_t> ecounts;
        allgatherv_u64(ctx, edges.p, 2 * (size_t)nE, alle, ecounts);
        size_t n2 = 0;
        for (size_t c : ecounts) n2 += c;
        VR_REQUIRE(n2 < 0xfffffff0ull, VRGPU_EUNSUPPORTED, "too many boundary edges");
        A.n_edges = n2 / 2;
        if (n2) {
            DBuf<uint64_t> keys(ctx, n2), ktmp(ctx, n2);
            DBuf<uint32_t> vals(ctx, n2), vtmp(ctx, n2), flags(ctx, n2), excl(ctx, n2), cidx(ctx, n2), tot(ctx, 1);
            VR_CUDA(cudaMemcpyAsync(keys.p, alle.p, n2 * 8, cudaMemcpyDeviceToDevice, ctx->stream));
            VR_LAUNCH(ctx, k_mg_iota, grid_for(n2, T), T, 0, vals.p, (uint32_t)n2);
            radix_sort_pairs(ctx, keys.p, vals.p, ktmp.p, vtmp.p, n2, 32 + bits_for_u64v((uint64_t)N));
            VR_LAUNCH(ctx, k_mg_heads, grid_for(n2, T), T, 0, keys.p, (uint32_t)n2, flags.p);
            exclusive_scan_u32(ctx, flags.p, excl.p, n2, tot.p);
            VR_CUDA(cudaMemcpyAsync(&nU, tot.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
            U.alloc(ctx, nU);
            VR_LAUNCH(ctx, k_mg_compact_ids, grid_for(n2, T), T, 0, keys.p, vals.p, flags.p, excl.p, (uint32_t)n2, cidx.p, U.p);
            DBuf<uint32_t> parent(ctx, nU);
            R.alloc(ctx, nU);
            VR_LAUNCH(ctx, k_uf_init, grid_for(nU, T), T, 0, parent.p, nU);
            VR_LAUNCH(ctx, k_mg_union, grid_for(n2 / 2, T), T, 0, cidx.p, (uint32_t)(n2 / 2), parent.p);
            VR_LAUNCH(ctx, k_uf_flatten, grid_for(nU, T), T, 0, parent.p, nU, R.p);
        }
    }
    A.n_boundary = nU;

    // ---- sizes and min pid: local per component, summed over ranks for the boundary components ----
    DBuf<uint32_t> csize(ctx, ncells + 1), bidx(ctx, ncells + 1);
    DBuf<unsigned long long> cminpid(ctx, ncells + 1);
    DBuf<unsigned long long> gsz(ctx, nU + 1), gmp(ctx, nU + 1);
    {
        KScope k(ctx, "mg.sizes");
        VR_CUDA(cudaMemsetAsync(csize.p, 0, sizeof(uint32_t) * (ncells + 1), ctx->stream));
        VR_CUDA(cudaMemsetAsync(bidx.p, 0xff, sizeof(uint32_t) * (ncells + 1), ctx->stream));
        VR_CUDA(cudaMemsetAsync(cminpid.p, 0xff, sizeof(unsigned long long) * (ncells + 1), ctx->stream));
        if (n_own) VR_LAUNCH(ctx, k_mg_comp_stats, grid_for(n_own, T), T, 0, comp.p, P.has_pid ? P.pid.p : (const long long *)nullptr, n_own, csize.p, cminpid.p);
        if (nU) {
            DBuf<unsigned long long> usize(ctx, nU), uminpid(ctx, nU);
            VR_LAUNCH(ctx, k_mg_boundary_contrib, grid_for(nU, T), T, 0, U.p, nU, (uint64_t)me, csize.p, cminpid.p, usize.p, uminpid.p, bidx.p);
            VR_NCCL(nccl_api().AllReduce(usize.p, usize.p, nU, ncclUint64, ncclSum, M.comm, ctx->stream));
            VR_NCCL(nccl_api().AllReduce(uminpid.p, uminpid.p, nU, ncclUint64, ncclMin, M.comm, ctx->stream));
            VR_CUDA(cudaMemsetAsync(gsz.p, 0, sizeof(unsigned long long) * nU, ctx->stream));
            VR_CUDA(cudaMemsetAsync(gmp.p, 0xff, sizeof(unsigned long long) * nU, ctx->stream));
            VR_LAUNCH(ctx, k_mg_boundary_totals, grid_for(nU, T), T, 0, R.p, nU, usize.p, uminpid.p, gsz.p, gmp.p);
        }
    }

    // ---- candidate groups -> global order ----
    const uint32_t ms = minsize < 1 ? 1u : (minsize > 0xffffffffLL ? 0xffffffffu : (uint32_t)minsize);
    GlobalOrder O;
    DBuf<uint32_t> gid_u(ctx, nU + 1), gid_cell(ctx, ncells + 1);
    {
        KScope k(ctx, "mg.global_order");
        DBuf<uint32_t> fl(ctx, ncells + 1), el(ctx, ncells + 1), fb(ctx, nU + 1), eb(ctx, nU + 1), tot(ctx, 2);
        VR_CUDA(cudaMemsetAsync(tot.p, 0, 2 * sizeof(uint32_t), ctx->stream));
        if (ncells) {
            VR_LAUNCH(ctx, k_mg_flag_local, grid_for(ncells, T), T, 0, S.root.p, bidx.p, csize.p, ncells, ms, fl.p);
            exclusive_scan_u32(ctx, fl.p, el.p, ncells, tot.p);
        }
        if (nU) {
            VR_LAUNCH(ctx, k_mg_flag_boundary, grid_for(nU, T), T, 0, R.p, U.p, nU, (uint64_t)me, gsz.p, ms, fb.p);
            exclusive_scan_u32(ctx, fb.p, eb.p, nU, tot.p + 1);
        }
        uint32_t h[2];
        VR_CUDA(cudaMemcpyAsync(h, tot.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        const size_t nmine = (size_t)h[0] + h[1];
        DBuf<uint64_t> rec(ctx, 3 * nmine + 1);
        if (h[0]) VR_LAUNCH(ctx, k_mg_records_local, grid_for(ncells, T), T, 0, fl.p, el.p, ncells, csize.p, cminpid.p, rec.p);
        if (h[1]) VR_LAUNCH(ctx, k_mg_records_boundary, grid_for(nU, T), T, 0, fb.p, eb.p, nU, gsz.p, gmp.p, rec.p + 3 * (size_t)h[0]);
        mg_global_order(ctx, rec.p, nmine, O);
        VR_CUDA(cudaMemsetAsync(gid_u.p, 0, sizeof(uint32_t) * (nU + 1), ctx->stream));
        VR_CUDA(cudaMemsetAsync(gid_cell.p, 0, sizeof(uint32_t) * (ncells + 1), ctx->stream));
        if (O.total) VR_LAUNCH(ctx, k_mg_scatter_ids, grid_for(O.total, T), T, 0, O.rec.p, O.gid_of.p, O.total, O.my0, O.my1, gid_u.p, gid_cell.p);
    }
    A.ngroups = O.total;
    A.pfof.alloc(ctx, n_own + 1);
    if (n_own)
        VR_LAUNCH(ctx, k_mg_write_pfof, grid_for(n_own, T), T, 0, comp.p, bidx.p, R.p, gid_u.p, gid_cell.p, n_own, A.pfof.p);
    // host copy of the global catalogue (size, home rank) for the placement of the per-group stages
    A.gsize.assign(O.total, 0);
    A.ghome.assign(O.total, -1);
    if (O.total) {
        std::vector<uint64_t> hrec(3 * (size_t)O.total);
        std::vector<uint32_t> hgid(O.total);
        VR_CUDA(cudaMemcpyAsync(hrec.data(), O.rec.p, hrec.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(hgid.data(), O.gid_of.p, hgid.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        int r = 0;
        for (uint32_t j = 0; j < O.total; j++) {
            while (j >= O.rank_first[r + 1]) r++;
            const uint32_t g = hgid[j];
            if (g == 0 || g > O.total) continue;
            A.gsize[g - 1] = (uint32_t)hrec[3 * (size_t)j];
            A.ghome[g - 1] = (hrec[3 * (size_t)j + 2] >> 63) ? -1 : r;
        }
    }
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
}

// Placement of the per-group stages (6DFOF, potentials, unbinding are independent per 3D group; the reference's
// MPIGroupExchange sends a group to the task holding most of it, mpiroutines.cxx:2548).  Here: groups that straddle a face
// or whose estimated cost matters go to the least loaded rank in size order (longest-processing-time rule); small groups
// held by one rank stay where they are and only enter as that rank's base load.  Every rank computes the same table.
void mg_place_groups(vrgpu_ctx *ctx, StageA &A, size_t n_own, const vrgpu_search_opts *o)
{
    MultiGpu &M = *ctx->mg;
    const int T = 256;
    const int N = M.nranks;
    const size_t G = A.gsize.size();
    const bool six = o->fofbgtype <= 4;
    const bool unbind = o->bound_halos >= 1;
    const long long pp = o->unbind.pp_threshold;
    auto cost = [&](double n) {
        double c = 0;
        if (six) c += 6.0 * n;                                                    // ns, measured orders of magnitude
        if (unbind) c += (pp < 0 || n <= (double)pp) ? 0.55e-3 * n * n : 0.45 * n * std::log2(n + 1.0) + 2.0 * n;
        return c;
    };
    const uint32_t BIG = 2048;
    std::vector<double> load(N, 0.0);
    for (size_t g = 0; g < G; g++)
        if (A.ghome[g] >= 0 && A.gsize[g] < BIG) load[A.ghome[g]] += cost(A.gsize[g]);
    std::vector<int> hdest(G + 1, -1);
    for (size_t g = 0; g < G; g++) { // global ids are in size order already
        int d = A.ghome[g];
        if (d < 0 || A.gsize[g] >= BIG) {
            d = (int)(std::min_element(load.begin(), load.end()) - load.begin());
            load[d] += cost(A.gsize[g]);
        }
        hdest[g + 1] = d;
    }
    DBuf<int> ddest(ctx, G + 1);
    VR_CUDA(cudaMemcpyAsync(ddest.p, hdest.data(), (G + 1) * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    A.dest.alloc(ctx, n_own + 1);
    if (n_own) VR_LAUNCH(ctx, k_mg_dest_of, grid_for(n_own, T), T, 0, A.pfof.p, ddest.p, n_own, A.dest.p);
    VR_CUDA(cudaStreamSynchronize(ctx->stream)); // hdest / ddest go out of scope
}

} // namespace
} // namespace vrgpu

namespace vrgpu {
namespace {

constexpr int MG_REC_WORDS = 9; // pos3, vel3, mass, pid, global group id

__global__ void k_mg_flag_dest(const int *__restrict__ dest, size_t n, int d, uint32_t *__restrict__ flags)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = dest[i] == d ? 1u : 0u;
}
__global__ void k_mg_pack_export(const double *__restrict__ pos, const double *__restrict__ vel, const double *__restrict__ mass,
                                 const long long *__restrict__ pid, const long long *__restrict__ pfof, const uint32_t *__restrict__ flags,
                                 const uint32_t *__restrict__ excl, size_t n, uint32_t *__restrict__ idx, uint64_t *__restrict__ out)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !flags[i]) return;
    const uint32_t k = excl[i];
    idx[k] = (uint32_t)i;
    double *o = reinterpret_cast<double *>(out + (size_t)MG_REC_WORDS * k);
    o[0] = pos[3 * i]; o[1] = pos[3 * i + 1]; o[2] = pos[3 * i + 2];
    o[3] = vel ? vel[3 * i] : 0.0; o[4] = vel ? vel[3 * i + 1] : 0.0; o[5] = vel ? vel[3 * i + 2] : 0.0;
    o[6] = mass ? mass[i] : 1.0;
    out[(size_t)MG_REC_WORDS * k + 7] = (uint64_t)(pid ? pid[i] : (long long)i);
    out[(size_t)MG_REC_WORDS * k + 8] = (uint64_t)pfof[i];
}
__global__ void k_mg_unpack_import(const uint64_t *__restrict__ in, size_t cnt, size_t at, double *__restrict__ pos, double *__restrict__ vel,
                                   double *__restrict__ mass, long long *__restrict__ pid, long long *__restrict__ pfof)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= cnt) return;
    const double *r = reinterpret_cast<const double *>(in + (size_t)MG_REC_WORDS * k);
    const size_t i = at + k;
    pos[3 * i] = r[0]; pos[3 * i + 1] = r[1]; pos[3 * i + 2] = r[2];
    vel[3 * i] = r[3]; vel[3 * i + 1] = r[4]; vel[3 * i + 2] = r[5];
    mass[i] = r[6];
    pid[i] = (long long)in[(size_t)MG_REC_WORDS * k + 7];
    pfof[i] = (long long)in[(size_t)MG_REC_WORDS * k + 8];
}
__global__ void k_mg_own_work_labels(const long long *__restrict__ pfof, const int *__restrict__ dest, size_t n, int me, long long *__restrict__ pfw)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pfw[i] = (dest[i] >= 0 && dest[i] != me) ? 0ll : pfof[i];
}
__global__ void k_mg_mark_ids(const long long *__restrict__ pfw, size_t n, uint32_t *__restrict__ present)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && pfw[i] > 0) present[pfw[i]] = 1u;
}
__global__ void k_mg_remap_ids(long long *__restrict__ pfw, size_t n, const uint32_t *__restrict__ present, const uint32_t *__restrict__ excl)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long g = pfw[i];
    if (g > 0) pfw[i] = (long long)excl[g] + 1; // dense local id, order of the global ids kept (size desc, min pid asc)
    (void)present;
}
__global__ void k_mg_final_stats(const long long *__restrict__ pfw, const long long *__restrict__ pid, size_t n, uint64_t *__restrict__ rec)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const long long g = i < n ? pfw[i] : 0ll;
    const bool on = g > 0;
    const unsigned peers = __match_any_sync(0xffffffffu, on ? (unsigned long long)g : 0ull); // one atomic per (warp, group)
    const unsigned long long mn = warp_min_u64(peers, on ? (unsigned long long)pid[i] : ~0ull);
    if (on && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) {
        atomicAdd((unsigned long long *)&rec[3 * (size_t)(g - 1)], (unsigned long long)__popc(peers));
        atomicMin((unsigned long long *)&rec[3 * (size_t)(g - 1) + 1], mn);
    }
}
__global__ void k_mg_init_final_rec(uint64_t *__restrict__ rec, uint32_t ng)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ng) return;
    rec[3 * (size_t)g] = 0; rec[3 * (size_t)g + 1] = ~0ull; rec[3 * (size_t)g + 2] = g;
}
__global__ void k_mg_to_global(long long *__restrict__ pfw, size_t n, const uint32_t *__restrict__ gid_of, uint32_t my0)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long g = pfw[i];
    pfw[i] = g > 0 ? (long long)gid_of[my0 + (uint32_t)(g - 1)] : 0ll;
}
__global__ void k_mg_scatter_back(const long long *__restrict__ lab, const uint32_t *__restrict__ idx, uint32_t cnt, long long *__restrict__ pfof)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < cnt) pfof[idx[k]] = lab[k];
}
// the same with the coordinates the consolidating rank rewrote: 7 words per particle (label, pos, vel)
__global__ void k_mg_pack_back(const long long *__restrict__ lab, const double *__restrict__ pos, const double *__restrict__ vel, size_t at, size_t cnt,
                               uint64_t *__restrict__ out)
{
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= cnt) return;
    const size_t i = at + k;
    out[7 * k] = (uint64_t)lab[i];
    double *o = reinterpret_cast<double *>(out + 7 * k + 1);
    o[0] = pos[3 * i]; o[1] = pos[3 * i + 1]; o[2] = pos[3 * i + 2];
    o[3] = vel[3 * i]; o[4] = vel[3 * i + 1]; o[5] = vel[3 * i + 2];
}
__global__ void k_mg_scatter_back_phase(const uint64_t *__restrict__ in, const uint32_t *__restrict__ idx, uint32_t cnt, long long *__restrict__ pfof,
                                        double *__restrict__ pos, double *__restrict__ vel)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= cnt) return;
    const size_t i = idx[k];
    pfof[i] = (long long)in[7 * (size_t)k];
    const double *r = reinterpret_cast<const double *>(in + 7 * (size_t)k + 1);
    pos[3 * i] = r[0]; pos[3 * i + 1] = r[1]; pos[3 * i + 2] = r[2];
    vel[3 * i] = r[3]; vel[3 * i + 1] = r[4]; vel[3 * i + 2] = r[5];
}

// ---- AdjustStructureForPeriod over all ranks, without moving a particle ------------------------------------------------
// The reference particle of a group is its member with the smallest pid (the first member of the reference's list whenever
// pids follow the input order; a rank does not know the caller's global order).  Its position reaches every rank through two
// all-reduces over per-group tables: MIN of the pid, then SUM of a table in which only the owner of that pid wrote its
// position (x + 0 is exact).  Every member is then shifted by whole periods exactly as search.cxx:871-887 does.
__global__ void k_mg_group_minpid(const long long *__restrict__ pfof, const long long *__restrict__ pid, size_t n, unsigned long long *__restrict__ minpid)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const long long g = i < n ? pfof[i] : 0ll;
    const bool on = g > 0;
    const unsigned peers = __match_any_sync(0xffffffffu, on ? (unsigned long long)g : 0ull); // one atomic per (warp, group)
    const unsigned long long mn = warp_min_u64(peers, on ? (unsigned long long)(pid ? pid[i] : (long long)i) : ~0ull);
    if (on && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) atomicMin(&minpid[g], mn);
}
__global__ void k_mg_ref_pos(const long long *__restrict__ pfof, const long long *__restrict__ pid, const double *__restrict__ pos, size_t n,
                             const unsigned long long *__restrict__ minpid, double *__restrict__ refpos)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long g = pfof[i];
    if (g > 0 && (unsigned long long)(pid ? pid[i] : (long long)i) == minpid[g]) {
        refpos[3 * (size_t)g] = pos[3 * i]; refpos[3 * (size_t)g + 1] = pos[3 * i + 1]; refpos[3 * (size_t)g + 2] = pos[3 * i + 2];
    }
}
__global__ void k_mg_adjust_ref(double *__restrict__ pos, const long long *__restrict__ pfof, size_t n, const double *__restrict__ refpos, double period,
                                uint32_t *__restrict__ moved_idx, double *__restrict__ moved_xyz, unsigned int *__restrict__ nmoved)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long g = pfof[i];
    if (g <= 0) return;
    double np[3];
    bool moved = false;
    for (int j = 0; j < 3; j++) {
        double x = pos[3 * i + j];
        const double diff = __dsub_rn(refpos[3 * (size_t)g + j], x);
        if (diff < -0.5 * period) { x = __dsub_rn(x, period); moved = true; }
        else if (diff > 0.5 * period) { x = __dadd_rn(x, period); moved = true; }
        np[j] = x;
    }
    if (moved) {
        const unsigned int slot = atomicAdd(nmoved, 1u);
        moved_idx[slot] = (uint32_t)i;
        for (int j = 0; j < 3; j++) { moved_xyz[3 * (size_t)slot + j] = np[j]; pos[3 * i + j] = np[j]; }
    }
}

// every rank calls this with the same (global) ngroups; returns the number of particles of P it moved
size_t mg_adjust_global(vrgpu_ctx *ctx, DeviceParticles &P, const long long *pfof, long long ngroups, double period, DBuf<uint32_t> &moved_idx,
                        DBuf<double> &moved_xyz)
{
    MultiGpu &M = *ctx->mg;
    const int T = 256;
    KScope k(ctx, "mg.adjust_for_period");
    const size_t ng1 = (size_t)ngroups + 1;
    DBuf<unsigned long long> minpid(ctx, ng1);
    DBuf<double> refpos(ctx, 3 * ng1);
    VR_CUDA(cudaMemsetAsync(minpid.p, 0xff, sizeof(unsigned long long) * ng1, ctx->stream));
    VR_CUDA(cudaMemsetAsync(refpos.p, 0, sizeof(double) * 3 * ng1, ctx->stream));
    const long long *pid = P.has_pid ? P.pid.p : nullptr;
    if (P.n) VR_LAUNCH(ctx, k_mg_group_minpid, grid_for(P.n, T), T, 0, pfof, pid, P.n, minpid.p);
    VR_NCCL(nccl_api().AllReduce(minpid.p, minpid.p, ng1, ncclUint64, ncclMin, M.comm, ctx->stream));
    if (P.n) VR_LAUNCH(ctx, k_mg_ref_pos, grid_for(P.n, T), T, 0, pfof, pid, P.pos.p, P.n, minpid.p, refpos.p);
    VR_NCCL(nccl_api().AllReduce(refpos.p, refpos.p, 3 * ng1, ncclDouble, ncclSum, M.comm, ctx->stream));
    moved_idx.alloc(ctx, P.n + 1);
    moved_xyz.alloc(ctx, 3 * P.n + 1);
    DBuf<unsigned int> nmoved(ctx, 1);
    VR_CUDA(cudaMemsetAsync(nmoved.p, 0, sizeof(unsigned int), ctx->stream));
    if (P.n) VR_LAUNCH(ctx, k_mg_adjust_ref, grid_for(P.n, T), T, 0, P.pos.p, pfof, P.n, refpos.p, period, moved_idx.p, moved_xyz.p, nmoved.p);
    unsigned int h = 0;
    VR_CUDA(cudaMemcpyAsync(&h, nmoved.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return h;
}

} // namespace

// SearchFullSet over all ranks.  P: this rank's slab (device resident, original order).  pfof_out: device array of P.n
// global group ids.  With wb the coordinates the path rewrites are rewritten in P as well (and wb says which, for the host
// write-back); without it P is only read.
void mg_search(vrgpu_ctx *ctx, DeviceParticles &P, double slab_lo, double slab_hi, const vrgpu_search_opts *o, DBuf<long long> &pfof_out,
               long long &ngroups, MgWriteback *wb)
{
    VR_REQUIRE(ctx->mg != nullptr, VRGPU_EINVAL, "vrgpu_dist_init has not been called");
    MultiGpu &M = *ctx->mg;
    const int T = 256;
    const int N = M.nranks, me = M.rank;
    const bool six = o->fofbgtype <= 4;
    const bool unbind = o->bound_halos >= 1;
    const size_t n_own = P.n;
    StageA A;
    mg_fof3d(ctx, P, slab_lo, slab_hi, std::sqrt(o->param1), o->period, o->halo_min_size, A);
    ctx->stats.n_groups3d = A.ngroups;
    if (wb && o->period > 0 && A.ngroups > 0) wb->nmoved = mg_adjust_global(ctx, P, A.pfof.p, A.ngroups, o->period, wb->midx, wb->mxyz);
    if (!(six || unbind) || A.ngroups == 0) {
        pfof_out = std::move(A.pfof);
        ngroups = A.ngroups;
        return;
    }
    VR_REQUIRE(P.has_vel, VRGPU_EINVAL, "6DFOF / unbinding need velocities");
    KScope kall(ctx, "mg.consolidate_total");
    // phase timers (each replaces the previous one: the scope ends where the next begins)
    std::unique_ptr<KScope> cphase(new KScope(ctx, "mg.c.place"));
    mg_place_groups(ctx, A, n_own, o);
    cphase.reset(); cphase.reset(new KScope(ctx, "mg.c.pack"));

    // ---- consolidate every boundary group on the rank of its root label (MPIGroupExchange) ----
    std::vector<DBuf<uint32_t>> eidx(N);
    std::vector<DBuf<uint64_t>> ebuf(N), ibuf(N);
    std::vector<unsigned long long> ecnt(N, 0);
    {
        DBuf<uint32_t> fl(ctx, n_own + 1), ex(ctx, n_own + 1), tot(ctx, 1);
        for (int d = 0; d < N; d++) {
            if (d == me || n_own == 0) continue;
            VR_LAUNCH(ctx, k_mg_flag_dest, grid_for(n_own, T), T, 0, A.dest.p, n_own, d, fl.p);
            exclusive_scan_u32(ctx, fl.p, ex.p, n_own, tot.p);
            uint32_t h = 0;
            VR_CUDA(cudaMemcpyAsync(&h, tot.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
            ecnt[d] = h;
            eidx[d].alloc(ctx, h + 1);
            ebuf[d].alloc(ctx, (size_t)MG_REC_WORDS * h + 1);
            if (h)
                VR_LAUNCH(ctx, k_mg_pack_export, grid_for(n_own, T), T, 0, P.pos.p, P.vel.p, P.has_mass ? P.mass.p : (const double *)nullptr,
                          P.has_pid ? P.pid.p : (const long long *)nullptr, A.pfof.p, fl.p, ex.p, n_own, eidx[d].p, ebuf[d].p);
        }
    }
    cphase.reset(); cphase.reset(new KScope(ctx, "mg.c.exchange"));
    std::vector<unsigned long long> allcnt = allgather_host<unsigned long long>(ctx, ecnt.data(), N); // allcnt[src*N + dst]
    size_t n_imp = 0;
    std::vector<size_t> ioff(N, 0);
    for (int s = 0; s < N; s++) { ioff[s] = n_imp; if (s != me) n_imp += (size_t)allcnt[(size_t)s * N + me]; }
    for (int s = 0; s < N; s++) if (s != me) ibuf[s].alloc(ctx, (size_t)MG_REC_WORDS * allcnt[(size_t)s * N + me] + 1);
    VR_NCCL(nccl_api().GroupStart());
    for (int d = 0; d < N; d++) {
        if (d == me) continue;
        if (ecnt[d]) VR_NCCL(nccl_api().Send(ebuf[d].p, (size_t)MG_REC_WORDS * 8 * ecnt[d], ncclUint8, d, M.comm, ctx->stream));
        const size_t rc = (size_t)allcnt[(size_t)d * N + me];
        if (rc) VR_NCCL(nccl_api().Recv(ibuf[d].p, (size_t)MG_REC_WORDS * 8 * rc, ncclUint8, d, M.comm, ctx->stream));
    }
    VR_NCCL(nccl_api().GroupEnd());

    cphase.reset(); cphase.reset(new KScope(ctx, "mg.c.workset"));
    // ---- working set: owned particles (exported ones unlabelled) + imported ----
    const size_t n_w = n_own + n_imp;
    DeviceParticles W;
    W.n = n_w;
    W.pos.alloc(ctx, 3 * n_w + 1); W.vel.alloc(ctx, 3 * n_w + 1); W.mass.alloc(ctx, n_w + 1); W.pid.alloc(ctx, n_w + 1);
    W.has_vel = W.has_mass = W.has_pid = true;
    DBuf<long long> pfw(ctx, n_w + 1);
    if (n_own) {
        VR_CUDA(cudaMemcpyAsync(W.pos.p, P.pos.p, 3 * n_own * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(W.vel.p, P.vel.p, 3 * n_own * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        if (P.has_mass) VR_CUDA(cudaMemcpyAsync(W.mass.p, P.mass.p, n_own * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        else { // (stream-ordered like everything else: the pool may hand out memory whose last use is still queued on ctx->stream)
            std::vector<double> one(n_own, 1.0);
            VR_CUDA(cudaMemcpyAsync(W.mass.p, one.data(), n_own * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
        }
        if (P.has_pid) VR_CUDA(cudaMemcpyAsync(W.pid.p, P.pid.p, n_own * sizeof(long long), cudaMemcpyDeviceToDevice, ctx->stream));
        else {
            std::vector<long long> id(n_own);
            std::iota(id.begin(), id.end(), 0ll);
            VR_CUDA(cudaMemcpyAsync(W.pid.p, id.data(), n_own * sizeof(long long), cudaMemcpyHostToDevice, ctx->stream));
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
        }
        VR_LAUNCH(ctx, k_mg_own_work_labels, grid_for(n_own, T), T, 0, A.pfof.p, A.dest.p, n_own, me, pfw.p);
    }
    for (int s = 0; s < N; s++) {
        if (s == me) continue;
        const size_t rc = (size_t)allcnt[(size_t)s * N + me];
        if (rc) VR_LAUNCH(ctx, k_mg_unpack_import, grid_for(rc, T), T, 0, ibuf[s].p, rc, n_own + ioff[s], W.pos.p, W.vel.p, W.mass.p, W.pid.p, pfw.p);
    }
    // dense local ids
    long long nloc = 0;
    {
        const size_t ngl = (size_t)A.ngroups + 1;
        DBuf<uint32_t> present(ctx, ngl), excl(ctx, ngl), tot(ctx, 1);
        VR_CUDA(cudaMemsetAsync(present.p, 0, sizeof(uint32_t) * ngl, ctx->stream));
        if (n_w) VR_LAUNCH(ctx, k_mg_mark_ids, grid_for(n_w, T), T, 0, pfw.p, n_w, present.p);
        exclusive_scan_u32(ctx, present.p, excl.p, ngl, tot.p);
        uint32_t h = 0;
        VR_CUDA(cudaMemcpyAsync(&h, tot.p, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        nloc = h;
        if (n_w) VR_LAUNCH(ctx, k_mg_remap_ids, grid_for(n_w, T), T, 0, pfw.p, n_w, present.p, excl.p);
    }

    cphase.reset();
    // ---- the single-GPU stages on the consolidated groups ----
    long long ngl = nloc;
    if (ngl > 0 && !wb) { // (with wb every group has been adjusted in P already, by the same rule)
        DBuf<uint32_t> midx;
        DBuf<double> mxyz;
        adjust_for_period_device(ctx, W, pfw.p, ngl, o->period, midx, mxyz, true);
    }
    if (six && ngl > 0) {
        DBuf<long long> pf6(ctx, n_w + 1);
        StageTimer t(ctx);
        const double param1 = o->param1 * (o->ellhalo6dxfac * o->ellhalo6dxfac);
        ngl = fof6d_device(ctx, W, pfw.p, ngl, std::sqrt(param1), o->ellhalo6dvfac, o->fofbgtype, o->halo_min_size, pf6.p, true);
        ctx->stats.fof6d_ms += t.stop();
        pfw = std::move(pf6);
    }
    if (unbind && ngl > 0) {
        int changed = 0;
        long long nmarg = 0;
        unbind_device(ctx, W, pfw.p, ngl, o->unbind, changed, nmarg);
        ctx->stats.n_marginal = nmarg;
    }

    cphase.reset(new KScope(ctx, "mg.c.renumber"));
    // ---- global renumbering of the surviving groups ----
    GlobalOrder O;
    {
        DBuf<uint64_t> rec(ctx, 3 * (size_t)ngl + 1);
        if (ngl) {
            VR_LAUNCH(ctx, k_mg_init_final_rec, grid_for((size_t)ngl, T), T, 0, rec.p, (uint32_t)ngl);
            VR_LAUNCH(ctx, k_mg_final_stats, grid_for(n_w, T), T, 0, pfw.p, W.pid.p, n_w, rec.p);
        }
        mg_global_order(ctx, rec.p, (size_t)ngl, O);
        if (n_w) VR_LAUNCH(ctx, k_mg_to_global, grid_for(n_w, T), T, 0, pfw.p, n_w, O.gid_of.p, O.my0);
    }
    ngroups = O.total;

    cphase.reset(); cphase.reset(new KScope(ctx, "mg.c.return"));
    // ---- labels of imported particles go home; after a 6DFOF, with wb, their rewritten coordinates as well ----
    pfof_out.alloc(ctx, n_own + 1);
    if (n_own) VR_CUDA(cudaMemcpyAsync(pfof_out.p, pfw.p, n_own * sizeof(long long), cudaMemcpyDeviceToDevice, ctx->stream));
    const bool phase = wb && six;
    if (phase && n_own) {
        // own particles that stayed here (the exported ones are unlabelled in W, so nothing touched their copies)
        VR_CUDA(cudaMemcpyAsync(P.pos.p, W.pos.p, 3 * n_own * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(P.vel.p, W.vel.p, 3 * n_own * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    const size_t bw = phase ? 7 : 1; // words per returned particle
    std::vector<DBuf<uint64_t>> back(N), bsend(N);
    for (int d = 0; d < N; d++) if (d != me) back[d].alloc(ctx, bw * ecnt[d] + 1);
    if (phase)
        for (int s = 0; s < N; s++) {
            const size_t rc = s == me ? 0 : (size_t)allcnt[(size_t)s * N + me];
            if (!rc) continue;
            bsend[s].alloc(ctx, 7 * rc);
            VR_LAUNCH(ctx, k_mg_pack_back, grid_for(rc, T), T, 0, pfw.p, W.pos.p, W.vel.p, n_own + ioff[s], rc, bsend[s].p);
        }
    VR_NCCL(nccl_api().GroupStart());
    for (int s = 0; s < N; s++) {
        if (s == me) continue;
        const size_t rc = (size_t)allcnt[(size_t)s * N + me];
        if (rc) VR_NCCL(nccl_api().Send(phase ? (const void *)bsend[s].p : (const void *)(pfw.p + n_own + ioff[s]), rc * 8 * bw, ncclUint8, s, M.comm, ctx->stream));
        if (ecnt[s]) VR_NCCL(nccl_api().Recv(back[s].p, (size_t)ecnt[s] * 8 * bw, ncclUint8, s, M.comm, ctx->stream));
    }
    VR_NCCL(nccl_api().GroupEnd());
    for (int d = 0; d < N; d++) {
        if (d == me || !ecnt[d]) continue;
        if (phase) VR_LAUNCH(ctx, k_mg_scatter_back_phase, grid_for((size_t)ecnt[d], T), T, 0, back[d].p, eidx[d].p, (uint32_t)ecnt[d], pfof_out.p, P.pos.p, P.vel.p);
        else VR_LAUNCH(ctx, k_mg_scatter_back, grid_for((size_t)ecnt[d], T), T, 0, reinterpret_cast<const long long *>(back[d].p), eidx[d].p, (uint32_t)ecnt[d], pfof_out.p);
    }
    if (wb) { wb->phase = phase; if (phase) wb->pf3 = std::move(A.pfof); }
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
}

} // namespace vrgpu
