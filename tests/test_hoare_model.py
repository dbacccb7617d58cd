"""CPU model for the one parity gap the design documents ("tree parity with tied coordinates"): the
reference's kd-tree picks the left half of a node with KDTree::MedianPos (NBodylib/src/KDTree/KDTree.cxx:304-337), a
quickselect whose Hoare partition decides by its swap history which of several particles with EXACTLY equal split
coordinate go left.  The device build sorts instead (ties -> lower member index), so node membership differs wherever a
median falls inside a run of equal coordinates -- common for float32-quantised snapshots.

This file is groundwork, not product code: it shows that one iteration of that partition is a DATA-PARALLEL operation
(two prefix counts, one pairing by rank, one fix-up), i.e. that the reference's arrangement can be reproduced exactly by
scan-based kernels, and pins the formulation against a literal transcription of MedianPos on inputs full of ties.

Parallel form of one iteration on positions [left, right] with pivot position k:
  x = X[p[k]]; swap p[k] <-> p[right];
  A = positions in [left, right) whose value is >= x, in increasing order  (the stops of the upward scan i)
  B = positions in [left, right) whose value is <= x, in decreasing order  (the stops of the downward scan j)
  M = number of m with A[m] < B[m]          (a prefix of the pairs, because A increases and B decreases)
  swap p[A[m]] <-> p[B[m]] for m < M
  i* = A[M] if A[M] exists and (M == 0 or A[M] < B[M-1]) else (B[M-1] if M > 0 else right)
  p[right], p[i*] = p[i*], pivot
  then right = i* - 1 if i* >= k, left = i* + 1 if i* <= k, as in the reference.
"""
import numpy as np
import pytest


def select_reference(perm, X, k, start, end):
    """Literal transcription of KDTree::MedianPos (balanced branch) on an index permutation, as oracle/vr_oracle.c does."""
    p = perm
    left, right = start, end - 1
    rounds = 0
    while left < right:
        rounds += 1
        x = X[p[k]]
        p[right], p[k] = p[k], p[right]
        i, j = left - 1, right
        while True:
            while i < j:
                i += 1
                if X[p[i]] >= x:
                    break
            while i < j:
                j -= 1
                if X[p[j]] <= x:
                    break
            w = p[i]
            p[i] = p[j]
            p[j] = w
            if j <= i:
                break
        p[j] = p[i]
        p[i] = p[right]
        p[right] = w
        if i >= k:
            right = i - 1
        if i <= k:
            left = i + 1
    return rounds


def select_parallel(perm, X, k, start, end):
    """The same result with every iteration written as whole-array operations (what scan kernels would do)."""
    p = perm
    left, right = start, end - 1
    rounds = 0
    while left < right:
        rounds += 1
        x = X[p[k]]
        p[right], p[k] = p[k], p[right]
        pos = np.arange(left, right)                       # the pivot now sits at `right` and is not scanned
        v = X[p[pos]]
        A = pos[v >= x]                                    # increasing
        B = pos[v <= x][::-1]                              # decreasing
        m = min(len(A), len(B))
        M = int(np.count_nonzero(A[:m] < B[:m]))           # a prefix: A increases, B decreases
        assert np.all(A[:M] < B[:M]) and (M == m or A[M] >= B[M])
        a, b = A[:M], B[:M]
        p[a], p[b] = p[b].copy(), p[a].copy()
        if M < len(A) and (M == 0 or A[M] < B[M - 1]):
            istar = int(A[M])
        elif M > 0:
            istar = int(B[M - 1])
        else:
            istar = right
        pivot = p[right]
        p[right] = p[istar]
        p[istar] = pivot
        if istar >= k:
            right = istar - 1
        if istar <= k:
            left = istar + 1
    return rounds


def _cases():
    rng = np.random.default_rng(11)
    for trial in range(400):
        n = int(rng.integers(2, 400))
        kind = trial % 5
        if kind == 0:
            X = rng.random(n)
        elif kind == 1:
            X = rng.integers(0, 4, size=n).astype(np.float64)              # almost everything ties
        elif kind == 2:
            X = np.round(rng.normal(size=n) * 3.0) / 2.0                    # runs of equal values around the median
        elif kind == 3:
            X = np.full(n, 7.0)                                             # all equal
        else:
            X = np.float32(50.0 + rng.normal(size=n) * 1e-5).astype(np.float64)  # float32-quantised coordinates
        start = int(rng.integers(0, max(1, n // 3)))
        end = int(rng.integers(start + 1, n + 1))
        k = start + (end - start - 1) // 2                                  # KDTree::BuildNodes' median index
        yield X, start, end, k, rng.permutation(n)


def test_parallel_partition_reproduces_quickselect_arrangement():
    total_rounds = 0
    for X, start, end, k, perm0 in _cases():
        p1, p2 = perm0.copy(), perm0.copy()
        r1 = select_reference(p1, X, k, start, end)
        r2 = select_parallel(p2, X, k, start, end)
        assert r1 == r2
        assert np.array_equal(p1, p2)                                       # the whole arrangement, not just the median
        # and it is a selection: everything left of k is <=, everything right of k is >=
        assert np.all(X[p1[start:k]] <= X[p1[k]]) and np.all(X[p1[k + 1:end]] >= X[p1[k]])
        total_rounds += r1
    assert total_rounds > 400


def test_recursive_build_matches_with_ties():
    """Whole trees: BuildNodes' recursion (max spread, first dimension wins, median index start + (size-1)/2, leaves of at
    most 8) with the two selections gives identical node membership and identical particle order on tied inputs."""
    rng = np.random.default_rng(5)

    def build(sel, pos, perm, start, end, out):
        size = end - start
        if size <= 8:
            return
        spread = [np.ptp(pos[perm[start:end], d]) for d in range(3)]
        d = int(np.argmax(spread))                                          # first maximum wins, like the reference
        k = start + (size - 1) // 2
        sel(perm, pos[:, d], k, start, end)
        out.append((start, k + 1, end, d))
        build(sel, pos, perm, start, k + 1, out)
        build(sel, pos, perm, k + 1, end, out)

    for n in (100, 1000, 3000):
        pos = np.float32(50.0 + rng.normal(size=(n, 3)) * 2e-5).astype(np.float64)  # heavy ties along every axis
        pa, pb = np.arange(n), np.arange(n)
        na, nb = [], []
        build(select_reference, pos, pa, 0, n, na)
        build(select_parallel, pos, pb, 0, n, nb)
        assert na == nb and np.array_equal(pa, pb)
        ties = sum(1 for (s, m, e, d) in na if pos[pa[m - 1], d] == pos[pa[m], d])
        if n >= 1000:
            assert ties > 0                                                 # medians do fall inside runs of equal coordinates


def select_device_form(perm, X, k, start, end, nt=8):
    """The iteration with the index arithmetic of the device kernels (csrc/ptree.cu, k_kd_big / ks_warp_select): both lists
    ASCENDING, B read from its end (B_desc[m] = B_asc[nB-1-m]), M by an nt-ary search for the first failing pair, hi exclusive."""
    p = perm
    lo, hi = start, end
    while lo + 1 < hi:
        x = X[p[k]]
        top = hi - 1
        p[top], p[k] = p[k], p[top]
        pos = np.arange(lo, top)
        v = X[p[pos]]
        An, Bn = pos[v >= x], pos[v <= x]
        nA, nB = len(An), len(Bn)
        mm = min(nA, nB)
        a, b = 0, mm
        while a < b:                                                        # k_kd_big's search
            step = (b - a + nt - 1) // nt
            first = nt
            for t in range(nt):
                m = a + t * step
                if m < b and not (An[m] < Bn[nB - 1 - m]):
                    first = t
                    break
            if first == nt:
                a = a + ((b - a - 1) // step) * step + 1
            else:
                b, a = a + first * step, (a + (first - 1) * step + 1 if first else a)
        M = a
        M2 = 0                                                              # ks_warp_select's chunked count with early exit
        for base in range(0, mm, 32):
            c = sum(1 for m in range(base, min(base + 32, mm)) if An[m] < Bn[nB - 1 - m])
            M2 += c
            if c < 32:
                break
        assert M == M2
        for m in range(M):
            ia, ib = An[m], Bn[nB - 1 - m]
            p[ia], p[ib] = p[ib], p[ia]
        if M < nA and (M == 0 or An[M] < Bn[nB - M]):
            istar = int(An[M])
        elif M > 0:
            istar = int(Bn[nB - M])
        else:
            istar = top
        p[istar], p[top] = p[top], p[istar]
        if istar >= k:
            hi = istar
        if istar <= k:
            lo = istar + 1


@pytest.mark.parametrize("nt", [2, 8, 1024])
def test_device_index_arithmetic_reproduces_quickselect(nt):
    for X, start, end, k, perm0 in _cases():
        p1, p2 = perm0.copy(), perm0.copy()
        select_reference(p1, X, k, start, end)
        select_device_form(p2, X, k, start, end, nt=nt)
        assert np.array_equal(p1, p2)
