// velociraptor-stf_b200/csrc/seqblock.cuh -- one CTA evaluates a SEQUENTIAL fp64 sum exactly (seqsum.h) -- shared by the per-group
// dispersion sums of the 6DFOF (fof6d.cu; search.cxx:443-493) and the per-node centre-of-mass sums of the kd-tree potentials
// (ptree.cu; unbind.cxx:1667-1676): both are plain left-to-right loops in the reference and the results that depend on them
// (6DFOF membership, which cells a tree walk opens) depend on their last bit.
#pragma once
#include "common.cuh"
#include "seqsum.h"

namespace vrgpu {

constexpr int SQ_NT = 512, SQ_PER = 8, SQ_W = SQ_NT * SQ_PER;
constexpr uint32_t SQ_MIN = 4096;

struct SeqShared {
    SeqMap wtot[SQ_NT / 32];
    double s;
    unsigned int firstbad;
};

// All SQ_NT threads of the CTA call this with the same arguments; term(k) is the k-th addend; every thread gets the sum.
template <class Term> __device__ double block_seqsum(Term term, unsigned long long n, double s0, SeqShared &sh)
{
    const unsigned tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    unsigned long long pos = 0;
    double s = s0;
    unsigned stretch = 32;
    while (pos < n) {
        long long S0;
        int eu;
        if (seq_decode_sum(s, S0, eu)) {
            const unsigned w = (unsigned)min((unsigned long long)SQ_W, n - pos);
            const unsigned k0 = tid * SQ_PER;
            SeqMap tm[SQ_PER];
            SeqMap inc = seq_identity();
#pragma unroll
            for (int j = 0; j < SQ_PER; j++) {
                tm[j] = (k0 + j < w) ? seq_term_map(term(pos + k0 + j), eu) : seq_identity();
                inc = seq_compose(inc, tm[j]);
            }
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                SeqMap o;
                o.e = __shfl_up_sync(0xffffffffu, inc.e, d);
                o.o = __shfl_up_sync(0xffffffffu, inc.o, d);
                if ((int)lane >= d) inc = seq_compose(o, inc);
            }
            if (lane == 31) sh.wtot[wid] = inc;
            if (tid == 0) sh.firstbad = w;
            __syncthreads();
            SeqMap pre = seq_identity();
            for (unsigned q = 0; q < wid; q++) pre = seq_compose(pre, sh.wtot[q]);
            SeqMap ex;
            ex.e = __shfl_up_sync(0xffffffffu, inc.e, 1);
            ex.o = __shfl_up_sync(0xffffffffu, inc.o, 1);
            if (lane == 0) ex = seq_identity();
            pre = seq_compose(pre, ex);
            long long S = seq_apply(S0, pre), Sb = 0;
            unsigned bad = w;
#pragma unroll
            for (int j = 0; j < SQ_PER; j++) {
                if (k0 + j < w && bad == w) {
                    const long long Sn = seq_apply(S, tm[j]);
                    if (!seq_in_binade(Sn)) { bad = k0 + j; Sb = S; }
                    else S = Sn;
                }
            }
            if (bad < w) atomicMin(&sh.firstbad, bad);
            __syncthreads();
            const unsigned fb = sh.firstbad;
            if (fb == w) {
                if (tid == (w - 1) / SQ_PER) sh.s = seq_encode_sum(S, eu);
            } else if (bad == fb)
                sh.s = fb > 0 ? seq_encode_sum(Sb, eu) : s;
            __syncthreads();
            s = sh.s;
            if (fb == w) {
                pos += w;
                stretch = 32;
                continue;
            }
            pos += fb;
            stretch = fb >= 64 ? 32u : min(stretch * 2u, 4096u);
        } else
            stretch = min(stretch * 2u, 4096u);
        // plain adds across the change of binade (or while the sum is zero / subnormal / not finite)
        const unsigned c = (unsigned)min((unsigned long long)stretch, n - pos);
        __syncthreads(); // everybody has read sh.s
        if (tid == 0) {
            double a = s;
            for (unsigned k = 0; k < c; k++) a = __dadd_rn(a, term(pos + k));
            sh.s = a;
        }
        __syncthreads();
        s = sh.s;
        pos += c;
    }
    __syncthreads();
    return s;
}

struct TermMomentum { // m v of member k (search.cxx:482-484)
    const double *v, *m;
    __device__ double operator()(unsigned long long k) const { return __dmul_rn(v[k], m[k]); }
};
struct TermMass {
    const double *m;
    __device__ double operator()(unsigned long long k) const { return m[k]; }
};
} // namespace vrgpu
