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

static int check(const char *name, const std::vector<double> &t, double s0 = 0.0)
{
    const double want = plain_sum(t, s0);
    int bad = 0;
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
        bad += check("squares*mass", t);
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
    printf(bad ? "seqsum: %d FAILURES\n" : "seqsum: all sequences bit-identical\n", bad);
    return bad ? 1 : 0;
}
