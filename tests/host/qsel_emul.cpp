// tests/host/qsel_emul.cpp -- serial emulation of the exact-ties kd-tree build (velociraptor-stf_b200/csrc/qsel.cuh).
//
// Test infrastructure.  The per-thread functions of qsel.cuh are compiled here by g++ and run once per "thread" in a
// chosen order (forward, reverse or shuffled: the result must not depend on the schedule inside a launch); the steps
// around them (frontier, min / max -> split dimension, the two scans of a round, children) are plain host restatements of
// the level loop in ptree.cu (potential_tree).  Input: several groups back to back (group g = positions
// [goff[g], goff[g+1])), each a root node.  Output: the final arrangement (tree position -> input index) and every
// node's [start, end), which tests/test_qsel_host.py compares with the oracle's BuildNodes restatement group by group.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <numeric>
#include <random>
#include <vector>

#include "qsel.cuh"

using namespace vrgpu::qsel;

namespace {

struct Order {
    std::vector<uint32_t> idx;
    Order(uint32_t n, int mode, std::mt19937_64 &g) : idx(n)
    {
        std::iota(idx.begin(), idx.end(), 0u);
        if (mode == 1) std::reverse(idx.begin(), idx.end());
        else if (mode == 2) std::shuffle(idx.begin(), idx.end(), g);
    }
};

} // namespace

// small_limit: nodes of at most this many particles take select_serial (QS_SMALL on the device; the tests also pass 0 and
// a few particles so that both paths see every size).  stats_out: [0] levels, [1] rounds, [2] nodes, [3] serial selections,
// [4] largest number of rounds in one level, [5] work items over all rounds.  Returns the number of nodes, or < 0 on error.
extern "C" long long qsel_emul(const double *pos, long long n, const long long *goff, long long ngroups, int bucket, int order_mode,
                               unsigned seed, unsigned small_limit, long long *perm_out, long long *node_out, long long node_cap,
                               long long *stats_out)
{
    for (int k = 0; k < 6; k++) stats_out[k] = 0;
    if (n <= 0 || ngroups <= 0) return 0;
    if (n >= 0x7fffffffll) return -1;
    std::mt19937_64 rng(seed);
    const uint32_t nL = (uint32_t)n;
    std::vector<double> x(nL), y(nL), z(nL);
    std::vector<uint32_t> slot_of(nL), perm(nL);
    for (uint32_t i = 0; i < nL; i++) { x[i] = pos[3 * i]; y[i] = pos[3 * i + 1]; z[i] = pos[3 * i + 2]; slot_of[i] = i; perm[i] = i; }

    std::vector<uint32_t> nstart, nend, nleft;
    std::vector<uint32_t> frontier;
    for (long long g = 0; g < ngroups; g++) {
        nstart.push_back((uint32_t)goff[g]);
        nend.push_back((uint32_t)goff[g + 1]);
        nleft.push_back(0xffffffffu);
        frontier.push_back((uint32_t)g);
    }
    std::vector<uint32_t> A(nL + 1), B(nL + 1), wf(nL + 1);
    std::vector<unsigned long long> flags(nL + 1), scan(nL + 1);

    for (int level = 0; level < 64; level++) {
        const uint32_t nF = (uint32_t)frontier.size();
        // min / max -> split flag and dimension (k_pt_minmax, k_pt_choose)
        std::vector<uint32_t> splitflag(nF), splitdim(nF);
        uint32_t nsplit = 0;
        for (uint32_t f = 0; f < nF; f++) {
            const uint32_t s = nstart[frontier[f]], e = nend[frontier[f]];
            splitflag[f] = e - s > (uint32_t)bucket ? 1u : 0u;
            splitdim[f] = 0;
            if (!splitflag[f]) continue;
            nsplit++;
            double best = 0;
            for (int d = 0; d < 3; d++) {
                const double *c = d == 0 ? x.data() : (d == 1 ? y.data() : z.data());
                double mn = c[perm[s]], mx = mn;
                for (uint32_t p = s; p < e; p++) { mn = std::min(mn, c[perm[p]]); mx = std::max(mx, c[perm[p]]); }
                const double sp = mx - mn;
                if (d == 0 || sp > best) { best = sp; splitdim[f] = (uint32_t)d; }
            }
        }
        if (nsplit == 0) break;
        stats_out[0]++;

        // ---- the selection (potential_tree, exact_ties branch) ----
        std::vector<uint32_t> lo(nF), hi(nF), npair(nF), len(nF), woff(nF);
        std::vector<double> pivot(nF);
        Level L{x.data(), y.data(), z.data(), slot_of.data(), perm.data(), frontier.data(), nstart.data(), nend.data(), splitflag.data(),
                splitdim.data(), lo.data(), hi.data(), npair.data(), len.data(), woff.data(), pivot.data(), wf.data(), flags.data(),
                scan.data(), A.data(), B.data(), nF, 0u};
        // init_thread uses the compile-time QS_SMALL; the emulation takes the limit as a parameter instead
        for (uint32_t f : Order(nF, order_mode, rng).idx) {
            const uint32_t s = nstart[frontier[f]], e = nend[frontier[f]];
            const bool big = splitflag[f] && e - s > small_limit;
            if (small_limit == QS_SMALL) {
                init_thread(L, f);
                if ((lo[f] != 0 || hi[f] != 0) != big) return -2;
            } else {
                lo[f] = big ? s : 0u;
                hi[f] = big ? e : 0u;
                len[f] = big && running(s, e) ? e - s - 1u : 0u;
            }
        }
        // k_qs_small
        for (uint32_t f : Order(nF, order_mode, rng).idx) {
            uint32_t s, e;
            if (small_limit == QS_SMALL ? !small_node(L, f, s, e)
                                        : !(s = nstart[frontier[f]], e = nend[frontier[f]], splitflag[f] && e - s <= small_limit))
                continue;
            std::vector<double> val(e - s);
            std::vector<uint32_t> idx(e - s);
            for (uint32_t i = 0; i < e - s; i++) { idx[i] = perm[s + i]; val[i] = coord(L, splitdim[f], idx[i]); }
            select_serial(val.data(), idx.data(), median_index(s, e) - s, e - s);
            for (uint32_t i = 0; i < e - s; i++) perm[s + i] = idx[i];
            stats_out[3]++;
        }
        long long rounds = 0;
        for (;;) {
            uint32_t run = 0;
            for (uint32_t f = 0; f < nF; f++) { woff[f] = run; run += len[f]; }
            L.W = run;
            if (L.W == 0) break;
            rounds++;
            stats_out[5] += L.W;
            for (uint32_t f : Order(nF, order_mode, rng).idx) begin_thread(L, f);
            for (uint32_t w : Order(L.W + 1, order_mode, rng).idx) flag_thread(L, w);
            unsigned long long acc = 0;
            for (uint32_t w = 0; w <= L.W; w++) { scan[w] = acc; acc += flags[w]; }
            for (uint32_t w : Order(L.W, order_mode, rng).idx) scatter_thread(L, w);
            for (uint32_t w : Order(L.W, order_mode, rng).idx) match_thread(L, w);
            for (uint32_t f : Order(nF, order_mode, rng).idx) end_thread(L, f);
            if (rounds > 100000) return -3;
        }
        stats_out[1] += rounds;
        stats_out[4] = std::max(stats_out[4], rounds);

        // children + next frontier (k_pt_children)
        std::vector<uint32_t> next;
        for (uint32_t f = 0; f < nF; f++) {
            const uint32_t nd = frontier[f];
            if (!splitflag[f]) { next.push_back(nd); continue; }
            const uint32_t s = nstart[nd], e = nend[nd], mid = median_index(s, e) + 1u;
            const uint32_t l = (uint32_t)nstart.size();
            nleft[nd] = l;
            nstart.push_back(s); nend.push_back(mid); nleft.push_back(0xffffffffu);
            nstart.push_back(mid); nend.push_back(e); nleft.push_back(0xffffffffu);
            next.push_back(l);
            next.push_back(l + 1);
        }
        frontier.swap(next);
    }
    for (uint32_t i = 0; i < nL; i++) perm_out[i] = perm[i];
    const long long nn = (long long)nstart.size();
    for (long long j = 0; j < nn && j < node_cap; j++) { node_out[2 * j] = nstart[j]; node_out[2 * j + 1] = nend[j]; }
    stats_out[2] = nn;
    return nn;
}
