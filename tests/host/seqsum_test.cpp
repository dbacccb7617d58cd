// Host unit test of velociraptor-stf_b200/csrc/seqsum.h: the windowed integer scan must reproduce the plain sequential fp64
// sum bit for bit.  The window loop below mirrors block_seqsum() in fof6d.cu with NT virtual threads of PER terms each.
#include "seqsum.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

using namespace vrgpu;

static long g_windows = 0, g_serial_terms = 0, g_scan_terms = 0;

static double windowed_sum(const std::vector<double> &t, double s0, int NT, int PER)
{
    const size_t n = t.size();
    const size_t W = (size_t)NT * PER;
    size_t pos = 0;
    double s = s0;
    size_t stretch = 32;
    std::vector<SeqMap> tmap(W), tot(NT), pfx(NT);
    while (pos < n) {
        long long S0;
        int eu;
        if (seq_decode_sum(s, S0, eu)) {
            const size_t w = std::min(W, n - pos);
            g_windows++;
            for (int th = 0; th < NT; th++) {
                SeqMap F = seq_identity();
                for (int j = 0; j < PER; j++) {
                    const size_t k = (size_t)th * PER + j;
                    tmap[k] = k < w ? seq_term_map(t[pos + k], eu) : seq_identity();
                    F = seq_compose(F, tmap[k]);
                }
                tot[th] = F;
            }
            SeqMap run = seq_identity();
            for (int th = 0; th < NT; th++) { pfx[th] = run; run = seq_compose(run, tot[th]); }
            size_t firstbad = w;
            long long Sbefore = S0, Slast = S0;
            for (int th = 0; th < NT && firstbad == w; th++) {
                long long S = seq_apply(S0, pfx[th]);
                for (int j = 0; j < PER; j++) {
                    const size_t k = (size_t)th * PER + j;
                    if (k >= w) break;
                    const long long Sn = seq_apply(S, tmap[k]);
                    if (!seq_in_binade(Sn)) { firstbad = k; Sbefore = S; break; }
                    S = Sn;
                    Slast = Sn;
                }
            }
            if (firstbad == w) {
                s = seq_encode_sum(Slast, eu);
                pos += w;
                g_scan_terms += w;
                stretch = 32;
                continue;
            }
            if (firstbad > 0) s = seq_encode_sum(Sbefore, eu);
            pos += firstbad;
            g_scan_terms += firstbad;
            stretch = firstbad >= 64 ? 32 : std::min<size_t>(stretch * 2, 4096);
        } else
            stretch = std::min<size_t>(stretch * 2, 4096);
        const size_t c = std::min(stretch, n - pos);
        for (size_t k = 0; k < c; k++) { volatile double v = s + t[pos + k]; s = v; }
        pos += c;
        g_serial_terms += c;
    }
    return s;
}

static double plain_sum(const std::vector<double> &t, double s0)
{
    double s = s0;
    for (double x : t) { volatile double v = s + x; s = v; }
    return s;
}

// The many-window evaluation of seqsum.h (mirrors k_seq_approx / k_seq_spec / k_seq_walk in fof6d.cu): summaries built
// under a predicted unit, one walker; windows whose summary does not apply take the windowed (always exact) path.
// `skew` perturbs the approximate prefix sums the prediction is taken from, `flip` corrupts every flip-th prediction.
static long g_fast = 0, g_slow = 0;
struct VecTerm {
    const std::vector<double> *t;
    double operator()(unsigned long long k) const { return (*t)[(size_t)k]; }
};
static double spec_sum(const std::vector<double> &t, double s0, size_t W, double skew, int flip)
{
    const size_t n = t.size(), nwin = (n + W - 1) / W;
    std::vector<double> start(nwin);
    {
        // inexact on purpose: per-window sums accumulated from the END of the window, prefix in long double
        long double run = s0;
        for (size_t k = 0; k < nwin; k++) {
            start[k] = (double)run * (1.0 + skew);
            double ws = 0;
            for (size_t j = std::min(n, (k + 1) * W); j-- > k * W;) ws += t[j];
            run += ws;
        }
    }
    std::vector<SeqSpec> spec(nwin);
    for (size_t k = 0; k < nwin; k++) {
        int eu = seq_predict_unit(start[k]);
        if (flip && k % (size_t)flip == (size_t)flip - 1 && eu != SEQ_NOSPEC) eu += (k & 1) ? 1 : -1;
        spec[k] = seq_spec_serial(VecTerm{&t}, k * W, (unsigned)std::min(W, n - k * W), eu);
    }
    double s = s0;
    for (size_t k = 0; k < nwin; k++) {
        long long S;
        int eu;
        if (seq_decode_sum(s, S, eu) && seq_spec_step(spec[k], S, eu)) {
            s = seq_encode_sum(S, eu);
            g_fast++;
        } else {
            std::vector<double> w(t.begin() + (long)(k * W), t.begin() + (long)std::min(n, (k + 1) * W));
            s = windowed_sum(w, s, 512, 8);
            g_slow++;
        }
    }
    return s;
}

// k_seq_spec's decomposition (NT virtual threads of PER terms: per-thread composite, exclusive prefix over threads, extreme
// prefix increments per thread, reduction) against the serial construction of seqsum.h
static SeqSpec spec_blocked(const std::vector<double> &t, size_t first, unsigned w, int eu, int NT, int PER)
{
    SeqSpec r{};
    r.eu = SEQ_NOSPEC;
    if (eu == SEQ_NOSPEC) return r;
    std::vector<SeqMap> tm((size_t)NT * PER), inc(NT);
    bool small = true;
    for (int th = 0; th < NT; th++) {
        inc[th] = seq_identity();
        for (int j = 0; j < PER; j++) {
            const size_t k = (size_t)th * PER + j;
            tm[k] = k < w ? seq_term_map(t[first + k], eu) : seq_identity();
            small = small && seq_term_small(tm[k]);
            inc[th] = seq_compose(inc[th], tm[k]);
        }
    }
    if (!small) return r;
    r.total = seq_identity();
    r.mine = r.mino = SEQ_HI;
    r.maxe = r.maxo = -SEQ_HI;
    SeqMap pre = seq_identity();
    for (int th = 0; th < NT; th++) {
        SeqMap p = pre;
        for (int j = 0; j < PER; j++) {
            const size_t k = (size_t)th * PER + j;
            if (k < w) {
                p = seq_compose(p, tm[k]);
                r.mine = std::min(r.mine, p.e); r.maxe = std::max(r.maxe, p.e);
                r.mino = std::min(r.mino, p.o); r.maxo = std::max(r.maxo, p.o);
            }
        }
        pre = seq_compose(pre, inc[th]);
    }
    r.total = pre;
    r.eu = eu;
    return r;
}
static int check_spec_shapes(const char *name, const std::vector<double> &t)
{
    int bad = 0;
    const size_t n = t.size();
    double run = 0;
    for (size_t first = 0; first < n; first += 4096) {
        const unsigned w = (unsigned)std::min<size_t>(4096, n - first);
        const int eu = seq_predict_unit(run);
        const SeqSpec a = seq_spec_serial(VecTerm{&t}, first, w, eu), b = spec_blocked(t, first, w, eu, 512, 8);
        if (a.eu != b.eu || (a.eu != SEQ_NOSPEC && (a.total.e != b.total.e || a.total.o != b.total.o || a.mine != b.mine || a.maxe != b.maxe ||
                                                     a.mino != b.mino || a.maxo != b.maxo))) {
            printf("FAIL %-28s summary of window %zu differs between the serial and the blocked construction\n", name, first / 4096);
            bad++;
            break;
        }
        for (unsigned k = 0; k < w; k++) run += t[first + k];
    }
    return bad;
}

static int check(const char *name, const std::vector<double> &t, double s0 = 0.0)
{
    const double want = plain_sum(t, s0);
    int bad = check_spec_shapes(name, t);
    {
        const long f0 = g_fast, sl0 = g_slow;
        const struct { size_t W; double skew; int flip; } sc[] = {{4096, 0.0, 0}, {4096, 1e-3, 0}, {1000, -0.3, 0}, {4096, 0.0, 3}, {37, 0.0, 0}};
        for (auto &c : sc) {
            const double got = spec_sum(t, s0, c.W, c.skew, c.flip);
            if (seq_bits(got) != seq_bits(want) && !(got != got && want != want)) {
                printf("FAIL %-28s many-window W=%zu skew=%g flip=%d n=%zu want %.17g got %.17g\n", name, c.W, c.skew, c.flip, t.size(), want, got);
                bad++;
            }
        }
        if (getenv("SEQSUM_VERBOSE")) printf("%-28s many-window: %ld fast, %ld slow windows\n", name, g_fast - f0, g_slow - sl0);
    }
    const long w0 = g_windows, sc0 = g_scan_terms, se0 = g_serial_terms;
    const int cfg[][2] = {{512, 8}, {32, 4}, {7, 3}, {1, 1}};
    for (auto &c : cfg) {
        const double got = windowed_sum(t, s0, c[0], c[1]);
        if (seq_bits(got) != seq_bits(want) && !(got != got && want != want)) {
            printf("FAIL %-28s NT=%d PER=%d n=%zu want %.17g (%016llx) got %.17g (%016llx)\n", name, c[0], c[1], t.size(), want,
                   (unsigned long long)seq_bits(want), got, (unsigned long long)seq_bits(got));
            bad++;
        }
    }
    if (getenv("SEQSUM_VERBOSE"))
        printf("%-28s n=%zu: windows %ld, scan terms %ld, serial terms %ld (all four shapes)\n", name, t.size(), g_windows - w0, g_scan_terms - sc0,
               g_serial_terms - se0);
    return bad;
}

int main()
{
    std::mt19937_64 rng(12345);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    std::normal_distribution<double> N(0.0, 1.0);
    int bad = 0;
    for (int rep = 0; rep < 6; rep++) {
        const size_t n = rep < 3 ? 100000 : 1000003;
        std::vector<double> t(n);
        // dispersion-like: m * (v - vbar)^2, float32-exact inputs
        for (auto &x : t) { const double v = (double)(float)(300.0 * N(rng)), m = (double)(float)0.0207; x = (v * v) * m; }
        {
            const long f0 = g_fast, sl0 = g_slow;
            bad += check("squares*mass", t);
            // exact prediction, W = 4096 is the first of the five shapes; over all five at least 60 % of the windows must
            // take the one-step path on a dispersion-like sum (the skewed / corrupted shapes lose theirs on purpose)
            if ((g_fast - f0) * 10 < (g_fast - f0 + g_slow - sl0) * 6) { printf("FAIL squares*mass: only %ld of %ld windows took the summary\n", g_fast - f0, g_fast - f0 + g_slow - sl0); bad++; }
        }
        // momentum-like: m * v with a drift
        for (auto &x : t) { const double v = (double)(float)(250.0 + 900.0 * N(rng)); x = v * 0.0207; }
        bad += check("drifting signed", t);
        // zero-mean random walk (sum crosses zero and many binades)
        for (auto &x : t) x = N(rng);
        bad += check("zero-mean walk", t);
        // equal masses (many exact ties: terms are multiples of a power of two)
        for (auto &x : t) x = 0.015625;
        bad += check("equal terms 2^-6", t);
        for (auto &x : t) x = 0.1;
        bad += check("equal terms 0.1", t, 1e-3);
        // forced ties: integers and half-integers against a sum near 2^53
        for (auto &x : t) x = (double)(rng() % 7) * 0.5;
        bad += check("half-integer ties", t, 4503599627370496.0 * 1.5);
        for (auto &x : t) x = ((rng() & 1) ? -1.0 : 1.0) * (double)(rng() % 9) * 0.5;
        bad += check("signed ties", t, 4503599627370496.0 * 1.25);
        // huge dynamic range, occasional giants
        for (auto &x : t) { x = std::ldexp(U(rng), (int)(rng() % 80) - 40); if ((rng() & 1023) == 0) x = -x * 1e6; }
        bad += check("wide range", t);
        // zeros, subnormals, sign flips
        for (auto &x : t) { const unsigned r = rng() & 7; x = r == 0 ? 0.0 : (r == 1 ? 4.9e-324 * (double)(rng() % 1000) : (r == 2 ? -1e-310 : N(rng) * 1e-300)); }
        bad += check("tiny / subnormal", t);
        // all zeros
        for (auto &x : t) x = 0.0;
        bad += check("all zero", t);
        // cancellation to exactly zero and restart
        for (size_t i = 0; i < n; i++) t[i] = (i & 1) ? -(double)((i >> 1) % 1000 + 1) : (double)((i >> 1) % 1000 + 1);
        bad += check("pairwise cancelling", t);
    }
    {
        std::vector<double> t(50000, 1.0);
        t[777] = INFINITY;
        bad += check("infinity", t);
        t[777] = NAN;
        bad += check("nan", t);
        t[777] = 1e308; t[778] = 1e308; t[779] = -1e308;
        bad += check("overflow", t);
    }
    printf("windows %ld, scan terms %ld, serial terms %ld (%.3f%%)\n", g_windows, g_scan_terms, g_serial_terms,
           100.0 * (double)g_serial_terms / (double)(g_serial_terms + g_scan_terms + 1));
    printf("many-window evaluation: %ld windows in one step, %ld the slow way\n", g_fast, g_slow);
    printf(bad ? "seqsum: %d FAILURES\n" : "seqsum: all sequences bit-identical\n", bad);
    return bad ? 1 : 0;
}
