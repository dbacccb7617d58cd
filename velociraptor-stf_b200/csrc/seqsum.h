// seqsum.h -- exact parallel evaluation of a SEQUENTIAL fp64 accumulation  s = fl(fl(fl(s0 + t0) + t1) + ...).
//
// Why: the reference computes each 3D group's velocity dispersion with plain left-to-right loops (search.cxx:443-493) and
// the 6DFOF membership depends on every bit of the result, so the GPU must reproduce the rounding of that order.  A
// dependent chain of DADDs costs ~50 ns per particle on one thread, i.e. seconds for a group of 10^7 members.
//
// How: while the running sum stays inside one binade [2^e, 2^(e+1)) its value is S*u with u = 2^(e-52) and S an integer in
// (2^52, 2^53), and fl(s + t) is "S + x rounded to the nearest integer, ties to even" with x = t/u.  Unless x is exactly
// half way between two integers the increment does not depend on S; if it is, it depends only on the parity of S.  So every
// term is a map  S -> S + (S even ? a_e : a_o), these maps are closed under composition, and composition is associative: a
// window of terms is folded with a parallel scan in integer arithmetic.  The scan result is accepted up to the first term
// whose result leaves the open interval (2^52, 2^53) -- there the binade (hence u) changes; a few terms are then added
// with real DADDs and the scan restarts in the new binade.  Results are bit-identical to the sequential loop for every
// input (zeros, mixed signs, cancellation, subnormals, infinities and NaNs take the plain-add path).
//
// This header is plain C++ shared by the CUDA kernels (fof6d.cu) and a host unit test (tests/host/seqsum_test.cpp).
#pragma once

#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define VR_HD __host__ __device__ __forceinline__
#else
#define VR_HD inline
#endif

namespace vrgpu {

struct SeqMap {
    long long e, o; // increment applied to an even / odd S
};

constexpr long long SEQ_LO = 1ll << 52, SEQ_HI = 1ll << 53;
constexpr long long SEQ_HUGE = 1ll << 58; // increment of a term that leaves the binade for sure (|t| >= 2^54 u, inf, nan)

VR_HD uint64_t seq_bits(double x)
{
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(x);
#else
    uint64_t b;
    memcpy(&b, &x, 8);
    return b;
#endif
}
VR_HD double seq_from_bits(uint64_t b)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)b);
#else
    double x;
    memcpy(&x, &b, 8);
    return x;
#endif
}

// s = S * 2^eu with |S| in [2^52, 2^53); false for zero, subnormal, infinite and NaN sums (plain adds handle those)
VR_HD bool seq_decode_sum(double s, long long &S, int &eu)
{
    const uint64_t b = seq_bits(s);
    const int ex = (int)((b >> 52) & 0x7ff);
    if (ex == 0 || ex == 0x7ff) return false;
    const long long m = (long long)((b & 0xfffffffffffffull) | (1ull << 52));
    S = (b >> 63) ? -m : m;
    eu = ex - 1075;
    return true;
}
// inverse of seq_decode_sum for |S| in [2^52, 2^53)
VR_HD double seq_encode_sum(long long S, int eu)
{
    const uint64_t m = (uint64_t)(S < 0 ? -S : S);
    return seq_from_bits((S < 0 ? (1ull << 63) : 0ull) | ((uint64_t)(eu + 1075) << 52) | (m & 0xfffffffffffffull));
}
VR_HD bool seq_in_binade(long long S)
{
    const uint64_t a = S < 0 ? (uint64_t)0 - (uint64_t)S : (uint64_t)S;
    return a > (uint64_t)SEQ_LO && a < (uint64_t)SEQ_HI;
}

// the map of one term for sums of unit 2^eu
VR_HD SeqMap seq_term_map(double t, int eu)
{
    const uint64_t b = seq_bits(t);
    const int ex = (int)((b >> 52) & 0x7ff);
    const bool neg = (b >> 63) != 0;
    uint64_t m = b & 0xfffffffffffffull;
    SeqMap r;
    if (ex == 0x7ff) { r.e = r.o = SEQ_HUGE; return r; } // inf / nan: force the plain-add path
    int et;
    if (ex == 0) {
        if (m == 0) { r.e = r.o = 0; return r; }
        et = -1074;
    } else {
        m |= 1ull << 52;
        et = ex - 1075;
    }
    const int shift = eu - et; // bits of m below the unit
    long long q;
    int cls; // 0 exact, 1 below half, 2 tie, 3 above half
    if (shift <= 0) {
        if (-shift > 1) { r.e = r.o = neg ? -SEQ_HUGE : SEQ_HUGE; return r; } // |t| >= 2^54 u: out of the binade for sure
        q = (long long)(m << (-shift));
        cls = 0;
    } else if (shift <= 63) {
        q = (long long)(m >> shift);
        const uint64_t rem = m & ((1ull << shift) - 1ull), half = 1ull << (shift - 1);
        cls = rem == 0 ? 0 : (rem < half ? 1 : (rem == half ? 2 : 3));
    } else {
        q = 0;
        cls = 1;
    }
    // x = t/u = +-(q + f); lowc = floor(x) when f > 0
    if (cls == 0) { r.e = r.o = neg ? -q : q; return r; }
    const long long lowc = neg ? -q - 1 : q;
    // f' = fractional part of x above lowc: f for positive x, 1 - f for negative x
    const int up = neg ? (cls == 1) : (cls == 3); // nearest integer is lowc + 1
    if (cls != 2) { r.e = r.o = lowc + (up ? 1 : 0); return r; }
    // tie: S + a must be even
    const long long ev = (lowc & 1) ? lowc + 1 : lowc, od = (lowc & 1) ? lowc : lowc + 1;
    r.e = ev;
    r.o = od;
    return r;
}

VR_HD SeqMap seq_identity()
{
    SeqMap r;
    r.e = r.o = 0;
    return r;
}
// wrapping add: composites that reach past the first out-of-binade term may overflow; they are never used
VR_HD long long seq_add(long long a, long long b) { return (long long)((uint64_t)a + (uint64_t)b); }
// first f, then h
VR_HD SeqMap seq_compose(const SeqMap &f, const SeqMap &h)
{
    SeqMap r;
    r.e = seq_add(f.e, (f.e & 1) ? h.o : h.e);
    r.o = seq_add(f.o, ((f.o + 1) & 1) ? h.o : h.e);
    return r;
}
VR_HD long long seq_apply(long long S, const SeqMap &a) { return seq_add(S, (S & 1) ? a.o : a.e); }

// ---------------------------------------------------------------------------------------------------------------------
// Many windows at once.  One CTA per group walks its windows one after the other (block_seqsum in fof6d.cu), which is a
// serial chain of ~n/4096 block scans: 0.3 s for a group of 2 10^7 members.  For such groups the windows are SUMMARISED
// in parallel instead, each under a PREDICTED unit 2^eu (taken from an ordinary, inexact parallel prefix sum of the
// terms), and one thread then walks the summaries:
//   * the summary holds the composed map of the window and, for an even and for an odd incoming S, the smallest and the
//     largest prefix increment inside the window;
//   * a window may be taken in one step iff the running sum really has the predicted unit and S + smallest, S + largest
//     both stay inside the binade with the sign of S (then every prefix does: the prefixes of one parity lie between their
//     extremes) -- exactly the condition under which block_seqsum would have accepted the whole window;
//   * otherwise (wrong prediction, change of binade inside the window, a term that is not small against the sum, sum
//     zero / subnormal / not finite) the window is evaluated the slow, always exact way.
// A window is summarised only if every increment is below 2^50 in magnitude: with at most 2^12 terms no prefix can then
// reach 2^62, so none of the integer additions can wrap.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int SEQ_NOSPEC = 0x7fffffff;      // SeqSpec::eu of a window without a usable summary
constexpr long long SEQ_SPEC_MAXINC = 1ll << 50;
constexpr unsigned SEQ_SPEC_MAXTERMS = 1u << 12;

struct SeqSpec {
    SeqMap total;                       // composed map of the whole window
    long long mine, maxe, mino, maxo;   // extreme prefix increments (over the non-empty prefixes) for an even / odd incoming S
    int eu;                             // unit exponent the summary was built for, or SEQ_NOSPEC
    int pad;
};

VR_HD bool seq_term_small(const SeqMap &m)
{
    return m.e > -SEQ_SPEC_MAXINC && m.e < SEQ_SPEC_MAXINC && m.o > -SEQ_SPEC_MAXINC && m.o < SEQ_SPEC_MAXINC;
}

// unit exponent predicted from an approximate value of the running sum (SEQ_NOSPEC if that value is not a normal number)
VR_HD int seq_predict_unit(double approx)
{
    long long S;
    int eu;
    return seq_decode_sum(approx, S, eu) ? eu : SEQ_NOSPEC;
}

// one step of the walk: true (and S advanced over the whole window) iff the summary applies to the sum S * 2^eu
VR_HD bool seq_spec_step(const SeqSpec &w, long long &S, int eu)
{
    if (w.eu == SEQ_NOSPEC || w.eu != eu) return false;
    const bool odd = (S & 1) != 0;
    const long long lo = S + (odd ? w.mino : w.mine), hi = S + (odd ? w.maxo : w.maxe); // |S| < 2^53, |increments| < 2^62
    if (!seq_in_binade(lo) || !seq_in_binade(hi) || ((lo < 0) != (S < 0)) || ((hi < 0) != (S < 0))) return false;
    S = seq_apply(S, w.total);
    return true;
}

// serial construction of a summary (host tests; the device builds the same numbers with a block scan)
template <class Term> VR_HD SeqSpec seq_spec_serial(Term term, unsigned long long first, unsigned count, int eu)
{
    SeqSpec w;
    w.total = seq_identity();
    w.mine = w.mino = SEQ_HI;
    w.maxe = w.maxo = -SEQ_HI;
    w.eu = eu;
    w.pad = 0;
    if (eu == SEQ_NOSPEC || count == 0 || count > SEQ_SPEC_MAXTERMS) { w.eu = SEQ_NOSPEC; return w; }
    for (unsigned k = 0; k < count; k++) {
        const SeqMap m = seq_term_map(term(first + k), eu);
        if (!seq_term_small(m)) { w.eu = SEQ_NOSPEC; return w; }
        w.total = seq_compose(w.total, m);
        w.mine = w.total.e < w.mine ? w.total.e : w.mine;
        w.maxe = w.total.e > w.maxe ? w.total.e : w.maxe;
        w.mino = w.total.o < w.mino ? w.total.o : w.mino;
        w.maxo = w.total.o > w.maxo ? w.total.o : w.maxo;
    }
    return w;
}

} // namespace vrgpu
