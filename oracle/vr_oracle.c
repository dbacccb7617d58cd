/* oracle/vr_oracle.c -- TEST INFRASTRUCTURE ONLY (see vr_oracle.h).
 *
 * Plain-C CPU restatement of the reference algorithms on the hot path.  Compile with -ffp-contract=off:
 * the reference default build emits no FMA and bit-exact membership depends on it.
 *
 * The FOF restatement does NOT re-create the reference's kd-tree BFS: FOF groups are the transitive closure of
 * the pair predicate, so any exact neighbour search gives the same partition provided every pair predicate is
 * evaluated with the reference's arithmetic (KDLeafNode.cxx:556-558, DistFunc.h:27-32).  The tree POTENTIAL does
 * re-create the reference's kd-tree (KDTree.cxx:438-507) because the monopole sum depends on the cells.
 */
#include "vr_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef long long i64;

#define NCMAX 320
#ifndef M_1_PI
#define M_1_PI 0.31830988618379067154 /* 1/pi, as <math.h> defines it outside strict C11 */
#endif

/* ------------------------------------------------------------------------------------------------ */
/* union-find                                                                                       */
/* ------------------------------------------------------------------------------------------------ */
static i64 uf_find(i64 *parent, i64 i)
{
    i64 r = i;
    while (parent[r] != r) r = parent[r];
    while (parent[i] != r) { i64 nx = parent[i]; parent[i] = r; i = nx; }
    return r;
}
static void uf_union(i64 *parent, i64 a, i64 b)
{
    a = uf_find(parent, a);
    b = uf_find(parent, b);
    if (a == b) return;
    if (a < b) parent[b] = a; else parent[a] = b;
}

/* DistanceSqd (DistFunc.h:27-32): three squared differences summed left to right, no FMA. */
static inline double dist2_3(const double *a, const double *b)
{
    return (a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]) + (a[2] - b[2]) * (a[2] - b[2]);
}

/* Pair predicate of LeafNode::FOFSearchBall (KDLeafNode.cxx:556-558): dist2 = DistanceSqd(pos);
 * if (numdim==6) dist2 += DistanceSqd(vel); link iff dist2 < fdist2 (strict).
 * `ti` is the target (possibly an image shifted by +-period, KDSplitNode.cxx:1500-1539). */
static inline int pair_links(const double *ti, const double *xj, const double *vi, const double *vj, double r2max)
{
    double d2 = dist2_3(ti, xj);
    if (vi) d2 += dist2_3(vi, vj);
    return d2 < r2max;
}

/* FOF6d (NBodylib/src/KDTree/FOFFunc.h:48-55), the comparison function of the 6D core search (search.cxx:1607-1611):
 * total += dx_j^2/params[6]; total += dv_j^2/params[7] for j = 0..2 in this order, link iff total < 1. */
static inline int pair_links_fof6d(const double *xi, const double *xj, const double *vi, const double *vj, const double *crit)
{
    double total = 0;
    for (int j = 0; j < 3; j++) {
        total += (xi[j] - xj[j]) * (xi[j] - xj[j]) / crit[0];
        total += (vi[j] - vj[j]) * (vi[j] - vj[j]) / crit[1];
    }
    return total < 1;
}

typedef struct { i64 key; i64 idx; } keyidx;
static int cmp_keyidx(const void *a, const void *b)
{
    const keyidx *x = (const keyidx *)a, *y = (const keyidx *)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx);
}

/* Exact neighbour search on a uniform grid over the first three coordinates + union-find.
 * x: n*3 positions, v: n*3 velocities or NULL, r2max: squared linking length, period: 0 for open. */
static void fof_link_crit(i64 n, const double *x, const double *v, double r2max, double period, const double *crit, i64 *parent);
static void fof_link(i64 n, const double *x, const double *v, double r2max, double period, i64 *parent)
{
    fof_link_crit(n, x, v, r2max, period, NULL, parent);
}
/* crit != NULL: the pair predicate is FOF6d with {params[6], params[7]} = crit (open box only); a linked pair has
 * |dx_j| < sqrt(params[6]) on every axis, which sizes the grid. */
static void fof_link_crit(i64 n, const double *x, const double *v, double r2max, double period, const double *crit, i64 *parent)
{
    for (i64 i = 0; i < n; i++) parent[i] = i;
    if (n < 2) return;
    if (crit) r2max = crit[0];
    double r = sqrt(r2max) * (1.0 + 1e-9);
    double lo[3], hi[3], cs[3];
    int nc[3];
    for (int k = 0; k < 3; k++) { lo[k] = hi[k] = x[k]; }
    for (i64 i = 0; i < n; i++)
        for (int k = 0; k < 3; k++) {
            if (x[3 * i + k] < lo[k]) lo[k] = x[3 * i + k];
            if (x[3 * i + k] > hi[k]) hi[k] = x[3 * i + k];
        }
    int brute = 0;
    for (int k = 0; k < 3; k++) {
        if (period > 0) {
            lo[k] = 0.0;
            double m = floor(period / r);
            if (m < 3) brute = 1;
            nc[k] = m > NCMAX ? NCMAX : (int)m;
            if (nc[k] < 1) nc[k] = 1;
            cs[k] = period / nc[k];
        } else {
            double ext = hi[k] - lo[k];
            double m = floor(ext / r);
            nc[k] = m > NCMAX ? NCMAX : (m < 1 ? 1 : (int)m);
            cs[k] = ext > 0 ? ext / nc[k] : 1.0;
            if (cs[k] < r) { nc[k] = 1; cs[k] = ext > 0 ? ext : 1.0; }
        }
    }
    /* keep the table small relative to n */
    while ((double)nc[0] * nc[1] * nc[2] > 8.0 * (double)n + 64 && !brute) {
        for (int k = 0; k < 3; k++) {
            if (nc[k] > 3) { nc[k] = (nc[k] + 1) / 2; if (nc[k] < 3) nc[k] = 3; }
            cs[k] = (period > 0 ? period : (hi[k] - lo[k] > 0 ? hi[k] - lo[k] : 1.0)) / nc[k];
        }
        if (nc[0] <= 3 && nc[1] <= 3 && nc[2] <= 3) break;
    }
    if (brute) {
        /* tiny periodic boxes (period < 3 r): test every pair against all 27 images, both directions */
        for (i64 i = 0; i < n; i++)
            for (i64 j = i + 1; j < n; j++) {
                int linked = 0;
                for (int sx = -1; sx <= 1 && !linked; sx++)
                    for (int sy = -1; sy <= 1 && !linked; sy++)
                        for (int sz = -1; sz <= 1 && !linked; sz++) {
                            double ti[3] = {x[3 * i] + sx * period, x[3 * i + 1] + sy * period, x[3 * i + 2] + sz * period};
                            double tj[3] = {x[3 * j] - sx * period, x[3 * j + 1] - sy * period, x[3 * j + 2] - sz * period};
                            if (pair_links(ti, &x[3 * j], v ? &v[3 * i] : NULL, v ? &v[3 * j] : NULL, r2max) ||
                                pair_links(tj, &x[3 * i], v ? &v[3 * j] : NULL, v ? &v[3 * i] : NULL, r2max))
                                linked = 1;
                        }
                if (linked) uf_union(parent, i, j);
            }
        return;
    }
    i64 ncell = (i64)nc[0] * nc[1] * nc[2];
    keyidx *ki = (keyidx *)malloc(sizeof(keyidx) * (size_t)n);
    int *cxyz = (int *)malloc(sizeof(int) * 3 * (size_t)n);
    for (i64 i = 0; i < n; i++) {
        int c[3];
        for (int k = 0; k < 3; k++) {
            double f = floor((x[3 * i + k] - lo[k]) / cs[k]);
            c[k] = f < 0 ? 0 : (f >= nc[k] ? nc[k] - 1 : (int)f);
            cxyz[3 * i + k] = c[k];
        }
        ki[i].key = ((i64)c[2] * nc[1] + c[1]) * nc[0] + c[0];
        ki[i].idx = i;
    }
    qsort(ki, (size_t)n, sizeof(keyidx), cmp_keyidx);
    i64 *cstart = (i64 *)malloc(sizeof(i64) * (size_t)(ncell + 1));
    {
        i64 p = 0;
        for (i64 c = 0; c <= ncell; c++) {
            while (p < n && ki[p].key < c) p++;
            cstart[c] = p;
        }
    }
    for (i64 a = 0; a < n; a++) {
        i64 i = ki[a].idx;
        const int *ci = &cxyz[3 * i];
        for (int dz = -1; dz <= 1; dz++)
            for (int dy = -1; dy <= 1; dy++)
                for (int dx = -1; dx <= 1; dx++) {
                    int c[3] = {ci[0] + dx, ci[1] + dy, ci[2] + dz};
                    double shift[3] = {0, 0, 0};
                    int ok = 1, wrapped = 0;
                    for (int k = 0; k < 3; k++) {
                        if (c[k] < 0) {
                            if (period > 0) { c[k] += nc[k]; shift[k] = period; wrapped = 1; } else ok = 0;
                        } else if (c[k] >= nc[k]) {
                            if (period > 0) { c[k] -= nc[k]; shift[k] = -period; wrapped = 1; } else ok = 0;
                        }
                    }
                    if (!ok) continue;
                    /* open box: out-of-range offsets are skipped above; periodic: nc >= 3 so offsets never alias */
                    i64 ck = ((i64)c[2] * nc[1] + c[1]) * nc[0] + c[0];
                    i64 me = ki[a].key;
                    if (ck < me) continue; /* each unordered cell pair once */
                    for (i64 bq = cstart[ck]; bq < cstart[ck + 1]; bq++) {
                        i64 j = ki[bq].idx;
                        if (ck == me && !wrapped && bq <= a) continue;
                        if (ck == me && wrapped && bq == a) continue;
                        if (uf_find(parent, i) == uf_find(parent, j)) continue;
                        int linked;
                        if (crit) {
                            linked = pair_links_fof6d(&x[3 * i], &x[3 * j], &v[3 * i], &v[3 * j], crit);
                        } else if (!wrapped) {
                            linked = pair_links(&x[3 * i], &x[3 * j], v ? &v[3 * i] : NULL, v ? &v[3 * j] : NULL, r2max);
                        } else {
                            /* target i imaged towards j, and target j imaged towards i (either BFS order can occur) */
                            double ti[3] = {x[3 * i] + shift[0], x[3 * i + 1] + shift[1], x[3 * i + 2] + shift[2]};
                            double tj[3] = {x[3 * j] - shift[0], x[3 * j + 1] - shift[1], x[3 * j + 2] - shift[2]};
                            linked = pair_links(ti, &x[3 * j], v ? &v[3 * i] : NULL, v ? &v[3 * j] : NULL, r2max) ||
                                     pair_links(tj, &x[3 * i], v ? &v[3 * j] : NULL, v ? &v[3 * i] : NULL, r2max);
                        }
                        if (linked) uf_union(parent, i, j);
                    }
                }
    }
    free(cstart);
    free(cxyz);
    free(ki);
}

/* Turn a parent forest into group ids: drop groups < minsize (KDFOF.cxx:74-84), order by size descending
 * (KDFOF.cxx:93-119) with the canonical min-pid tie break.  Returns number of groups. */
typedef struct { i64 size; i64 minpid; i64 root; } grec;
static int cmp_grec(const void *a, const void *b)
{
    const grec *x = (const grec *)a, *y = (const grec *)b;
    if (x->size != y->size) return x->size > y->size ? -1 : 1;
    if (x->minpid != y->minpid) return x->minpid < y->minpid ? -1 : 1;
    return 0;
}
static i64 label_groups(i64 n, i64 *parent, i64 minsize, const i64 *pid, i64 *pfof_out, i64 id_offset)
{
    i64 *cnt = (i64 *)calloc((size_t)n + 1, sizeof(i64));
    i64 *mp = (i64 *)malloc(sizeof(i64) * ((size_t)n + 1));
    for (i64 i = 0; i < n; i++) mp[i] = 0x7fffffffffffffffLL;
    for (i64 i = 0; i < n; i++) {
        i64 r = uf_find(parent, i);
        cnt[r]++;
        i64 p = pid ? pid[i] : i;
        if (p < mp[r]) mp[r] = p;
    }
    i64 ng = 0;
    for (i64 i = 0; i < n; i++) if (cnt[i] >= minsize && cnt[i] > 0 && parent[i] == i) ng++;
    grec *g = (grec *)malloc(sizeof(grec) * (size_t)(ng + 1));
    i64 q = 0;
    for (i64 i = 0; i < n; i++)
        if (cnt[i] >= minsize && cnt[i] > 0 && parent[i] == i) { g[q].size = cnt[i]; g[q].minpid = mp[i]; g[q].root = i; q++; }
    qsort(g, (size_t)ng, sizeof(grec), cmp_grec);
    for (i64 i = 0; i < n; i++) cnt[i] = 0; /* reuse as root -> id */
    for (i64 k = 0; k < ng; k++) cnt[g[k].root] = k + 1 + id_offset;
    for (i64 i = 0; i < n; i++) pfof_out[i] = cnt[uf_find(parent, i)];
    free(g);
    free(mp);
    free(cnt);
    return ng;
}

long long vro_fof3d(long long n, const double *pos, double b, const double *period, long long minsize,
                    const long long *pid, long long *pfof_out)
{
    if (n <= 0) return 0;
    i64 *parent = (i64 *)malloc(sizeof(i64) * (size_t)n);
    /* KDFOF.cxx:14: fdist2 = fdist*fdist */
    fof_link(n, pos, NULL, b * b, period ? period[0] : 0.0, parent);
    i64 ng = label_groups(n, parent, minsize, pid, pfof_out, 0);
    free(parent);
    return ng;
}

/* KDTree::FOFCriterion(FOF6d, params, numgroups, minsize, order = 1, ipcheckflag = 1, FOFcheckbg) as the 6D core search calls
 * it (search.cxx:1607-1611, 1664; KDFOF.cxx:133-240): particles whose potential field is >= params[9] are left out
 * (FOFcheckbg, fofalgo.cxx:205-208), the others are linked by FOF6d; groups below minsize dissolve; ids by size descending
 * with the canonical min-pid tie break.  The tree only prunes with a lower bound of the position part of `total`
 * (KDSplitNode.cxx:894-925 with params[1] == params[6]), so the groups are the connected components of the predicate. */
long long vro_fof_criterion6d(long long n, const double *pos, const double *vel, const double *pot, double p6, double p7,
                              double p9, long long minsize, const long long *pid, long long *pfof_out)
{
    if (n <= 0) return 0;
    i64 na = 0;
    i64 *act = (i64 *)malloc(sizeof(i64) * (size_t)n);
    for (i64 i = 0; i < n; i++) { pfof_out[i] = 0; if (pot[i] < p9) act[na++] = i; }
    if (na == 0) { free(act); return 0; }
    double *x = (double *)malloc(sizeof(double) * 3 * (size_t)na), *v = (double *)malloc(sizeof(double) * 3 * (size_t)na);
    i64 *cp = (i64 *)malloc(sizeof(i64) * (size_t)na), *parent = (i64 *)malloc(sizeof(i64) * (size_t)na);
    i64 *out = (i64 *)malloc(sizeof(i64) * (size_t)na);
    for (i64 a = 0; a < na; a++) {
        for (int k = 0; k < 3; k++) { x[3 * a + k] = pos[3 * act[a] + k]; v[3 * a + k] = vel[3 * act[a] + k]; }
        cp[a] = pid ? pid[act[a]] : act[a];
    }
    const double crit[2] = {p6, p7};
    fof_link_crit(na, x, v, 0.0, 0.0, crit, parent);
    i64 ng = label_groups(na, parent, minsize, cp, out, 0);
    for (i64 a = 0; a < na; a++) pfof_out[act[a]] = out[a];
    free(out); free(parent); free(cp); free(v); free(x); free(act);
    return ng;
}

void vro_adjust_for_period(long long n, double *pos, double period, const long long *pfof, long long ngroups)
{
    /* search.cxx:865-891: reference = first member in index order; others moved to its nearest image */
    i64 *iref = (i64 *)malloc(sizeof(i64) * (size_t)(ngroups + 1));
    for (i64 g = 0; g <= ngroups; g++) iref[g] = -1;
    for (i64 i = 0; i < n; i++) {
        if (pfof[i] == 0) continue;
        if (iref[pfof[i]] != -1) continue;
        iref[pfof[i]] = i;
    }
    for (i64 i = 0; i < n; i++) {
        i64 g = pfof[i];
        if (g == 0 || iref[g] == i) continue;
        for (int j = 0; j < 3; j++) {
            double diff = pos[3 * iref[g] + j] - pos[3 * i + j];
            if (diff < -0.5 * period) pos[3 * i + j] = pos[3 * i + j] - period;
            else if (diff > 0.5 * period) pos[3 * i + j] = pos[3 * i + j] + period;
        }
    }
    free(iref);
}

long long vro_fof6d(long long n, double *pos, double *vel, const double *mass, const long long *pfof3d,
                    long long ngroups3d, double b6, double vfac, int mode, long long minsize,
                    const long long *pid, long long *pfof_out, double *vscale2_out)
{
    for (i64 i = 0; i < n; i++) pfof_out[i] = 0;
    if (ngroups3d <= 0) return 0;
    /* search.cxx:422-433: counts and offsets; members gathered in index order */
    i64 *num = (i64 *)calloc((size_t)ngroups3d + 2, sizeof(i64));
    i64 *off = (i64 *)calloc((size_t)ngroups3d + 2, sizeof(i64));
    for (i64 i = 0; i < n; i++) if (pfof3d[i] > 0) num[pfof3d[i]]++;
    off[1] = 0;
    for (i64 g = 2; g <= ngroups3d + 1; g++) off[g] = off[g - 1] + num[g - 1];
    i64 ntot = off[ngroups3d + 1];
    i64 *memb = (i64 *)malloc(sizeof(i64) * (size_t)(ntot + 1));
    {
        i64 *fill = (i64 *)calloc((size_t)ngroups3d + 2, sizeof(i64));
        for (i64 i = 0; i < n; i++) if (pfof3d[i] > 0) { i64 g = pfof3d[i]; memb[off[g] + fill[g]++] = i; }
        free(fill);
    }
    double *vs2 = (double *)calloc((size_t)ngroups3d + 2, sizeof(double));
    if (mode == 4) {
        /* search.cxx:443-470 including quirk Q1: mtotregion is incremented ONCE, after the loop, with the
         * particle that follows the largest group in the group-sorted array */
        double vx = 0, vy = 0, vz = 0, mtot = 0, vscale2 = 0;
        for (i64 a = 0; a < num[1]; a++) {
            i64 i = memb[off[1] + a];
            vx += vel[3 * i] * mass[i]; vy += vel[3 * i + 1] * mass[i]; vz += vel[3 * i + 2] * mass[i];
        }
        i64 inext = -1;
        if (ngroups3d >= 2) inext = memb[off[2]];
        else for (i64 i = 0; i < n; i++) if (pfof3d[i] == 0) { inext = i; break; }
        mtot += inext >= 0 ? mass[inext] : mass[0];
        double vm[3] = {vx / mtot, vy / mtot, vz / mtot};
        for (i64 a = 0; a < num[1]; a++) {
            i64 i = memb[off[1] + a];
            for (int k = 0; k < 3; k++) vscale2 += pow(vel[3 * i + k] - vm[k], 2.0) * mass[i];
        }
        if (mtot > 0) vscale2 /= mtot;
        vscale2 *= vfac * vfac;
        for (i64 g = 1; g <= ngroups3d; g++) vs2[g] = vscale2;
    } else {
        /* search.cxx:480-493 */
        for (i64 g = 1; g <= ngroups3d; g++) {
            double vx = 0, vy = 0, vz = 0, mtot = 0, vscale2 = 0;
            for (i64 a = 0; a < num[g]; a++) {
                i64 i = memb[off[g] + a];
                vx += vel[3 * i] * mass[i]; vy += vel[3 * i + 1] * mass[i]; vz += vel[3 * i + 2] * mass[i];
                mtot += mass[i];
            }
            double vm[3] = {vx / mtot, vy / mtot, vz / mtot};
            for (i64 a = 0; a < num[g]; a++) {
                i64 i = memb[off[g] + a];
                for (int k = 0; k < 3; k++) vscale2 += pow(vel[3 * i + k] - vm[k], 2.0) * mass[i];
            }
            vs2[g] = vscale2 / mtot * vfac * vfac;
        }
    }
    if (vscale2_out) for (i64 g = 0; g <= ngroups3d; g++) vscale2_out[g] = vs2[g];

    /* search.cxx:528-535: in mode 4 all groups with n <= ompunbindnum(1000) form ONE search chunk */
    i64 iend = ngroups3d;
    if (mode == 4) for (i64 g = 1; g <= ngroups3d; g++) if (num[g] <= 1000) { iend = g; break; }

    i64 ng = 0;
    i64 nmaxchunk = 0;
    for (i64 g = 1; g <= iend; g++) {
        i64 cnt = (g == iend && mode == 4) ? (ntot - off[g]) : num[g];
        if (cnt > nmaxchunk) nmaxchunk = cnt;
    }
    double *x = (double *)malloc(sizeof(double) * 3 * (size_t)(nmaxchunk + 1));
    double *v = (double *)malloc(sizeof(double) * 3 * (size_t)(nmaxchunk + 1));
    i64 *parent = (i64 *)malloc(sizeof(i64) * (size_t)(nmaxchunk + 1));
    i64 *lp = (i64 *)malloc(sizeof(i64) * (size_t)(nmaxchunk + 1));
    i64 *lpid = (i64 *)malloc(sizeof(i64) * (size_t)(nmaxchunk + 1));
    for (i64 g = 1; g <= iend; g++) {
        i64 cnt = (g == iend && mode == 4) ? (ntot - off[g]) : num[g];
        if (cnt <= 0) continue;
        /* search.cxx:565-568: scale in place (one rounding per coordinate) */
        double xscaling = 1.0 / b6, vscaling = 1.0 / sqrt(vs2[g]);
        for (i64 a = 0; a < cnt; a++) {
            i64 i = memb[off[g] + a];
            for (int k = 0; k < 3; k++) { pos[3 * i + k] *= xscaling; vel[3 * i + k] *= vscaling; }
            for (int k = 0; k < 3; k++) { x[3 * a + k] = pos[3 * i + k]; v[3 * a + k] = vel[3 * i + k]; }
            lpid[a] = pid ? pid[i] : i;
        }
        xscaling = 1.0 / xscaling; vscaling = 1.0 / vscaling;
        /* search.cxx:570-571: 6-D tree, FOF(1.0) non periodic */
        fof_link(cnt, x, v, 1.0 * 1.0, 0.0, parent);
        i64 ngl = label_groups(cnt, parent, minsize, lpid, lp, 0);
        /* search.cxx:573-575 */
        for (i64 a = 0; a < cnt; a++) {
            i64 i = memb[off[g] + a];
            for (int k = 0; k < 3; k++) { pos[3 * i + k] *= xscaling; vel[3 * i + k] *= vscaling; }
            /* search.cxx:664 */
            pfof_out[i] = lp[a] + (lp[a] > 0) * ng;
        }
        ng += ngl;
    }
    free(lpid); free(lp); free(parent); free(v); free(x);
    free(vs2); free(memb); free(off); free(num);
    /* search.cxx:681-689: ReorderGroupIDs by size -> canonical (size desc, min pid asc) */
    if (ng > 0) {
        grec *gr = (grec *)malloc(sizeof(grec) * (size_t)(ng + 1));
        for (i64 k = 0; k <= ng; k++) { gr[k].size = 0; gr[k].minpid = 0x7fffffffffffffffLL; gr[k].root = k; }
        for (i64 i = 0; i < n; i++) if (pfof_out[i] > 0) {
            grec *r = &gr[pfof_out[i]];
            r->size++;
            i64 p = pid ? pid[i] : i;
            if (p < r->minpid) r->minpid = p;
        }
        qsort(gr + 1, (size_t)ng, sizeof(grec), cmp_grec);
        i64 *newid = (i64 *)calloc((size_t)ng + 1, sizeof(i64));
        for (i64 k = 1; k <= ng; k++) newid[gr[k].root] = k;
        for (i64 i = 0; i < n; i++) if (pfof_out[i] > 0) pfof_out[i] = newid[pfof_out[i]];
        free(newid);
        free(gr);
    }
    return ng;
}

/* ------------------------------------------------------------------------------------------------ */
/* potentials                                                                                       */
/* ------------------------------------------------------------------------------------------------ */
void vro_potential_pp(long long n, const double *pos, const double *mass, double G, double eps, double *phi)
{
    /* unbind.cxx:402-411 */
    double eps2 = eps * eps;
    for (i64 j = 0; j < n; j++) phi[j] = 0;
    for (i64 j = 0; j < n; j++)
        for (i64 k = j + 1; k < n; k++) {
            double r2 = 0.;
            for (int d = 0; d < 3; d++) r2 += pow(pos[3 * j + d] - pos[3 * k + d], 2.0);
            r2 += eps2;
            r2 = 1.0 / sqrt(r2);
            double pot = -G * (mass[j] * mass[k]) * r2;
            phi[j] = phi[j] + pot;
            phi[k] = phi[k] + pot;
        }
}

typedef struct {
    i64 start, end;      /* particle range in tree order */
    i64 left, right;     /* child node ids, -1 for leaves */
    double cm[3], mtot, r2max;
} tnode;

typedef struct {
    const double *pos;   /* original positions */
    i64 *perm;           /* tree order -> original index */
    int bucket;
    tnode *nodes;
    i64 nnodes, cap;
} tbuild;

/* KDTree::MedianPos (KDTree.cxx:304-337) restated on an index permutation: after the call perm[k] holds the
 * element of rank k along dimension d within [start,end), smaller-or-equal ones to its left. */
static void select_median(tbuild *t, int d, i64 k, i64 start, i64 end)
{
    i64 left = start, right = end - 1;
    i64 *p = t->perm;
    const double *X = t->pos;
    while (left < right) {
        double x = X[3 * p[k] + d];
        i64 w = p[right]; p[right] = p[k]; p[k] = w;
        i64 i = left - 1, j = right;
        for (;;) {
            while (i < j) if (X[3 * p[++i] + d] >= x) break;
            while (i < j) if (X[3 * p[--j] + d] <= x) break;
            w = p[i]; p[i] = p[j]; p[j] = w;
            if (j <= i) break;
        }
        p[j] = p[i];
        p[i] = p[right];
        p[right] = w;
        if (i >= k) right = i - 1;
        if (i <= k) left = i + 1;
    }
}

/* KDTree::BuildNodes (KDTree.cxx:438-507) with the default splitting criterion (max spread, first wins),
 * nodes numbered in the pre-order of GetNodeList (unbind.cxx:13-21). */
static i64 build_nodes(tbuild *t, i64 start, i64 end)
{
    if (t->nnodes == t->cap) {
        t->cap = t->cap * 2 + 16;
        t->nodes = (tnode *)realloc(t->nodes, sizeof(tnode) * (size_t)t->cap);
    }
    i64 id = t->nnodes++;
    t->nodes[id].start = start;
    t->nodes[id].end = end;
    t->nodes[id].left = t->nodes[id].right = -1;
    i64 size = end - start;
    if (size <= t->bucket) return id;
    int splitdim = 0;
    double maxspread = 0;
    for (int j = 0; j < 3; j++) {
        double mn = t->pos[3 * t->perm[start] + j], mx = mn;
        for (i64 i = start + 1; i < end; i++) {
            double c = t->pos[3 * t->perm[i] + j];
            if (c < mn) mn = c;
            if (c > mx) mx = c;
        }
        double spread = mx - mn;
        if (j == 0 || spread > maxspread) { if (j == 0) { maxspread = spread; splitdim = 0; } else { maxspread = spread; splitdim = j; } }
    }
    i64 k = start + (size - 1) / 2;
    select_median(t, splitdim, k, start, end);
    i64 l = build_nodes(t, start, k + 1);
    i64 r = build_nodes(t, k + 1, end);
    t->nodes[id].left = l;
    t->nodes[id].right = r;
    return id;
}

typedef struct {
    const tbuild *t;
    const double *tp;    /* positions in tree order */
    const double *tm;    /* masses in tree order */
    double eps2;
    double *r2val;       /* accepted-cell terms M/sqrt(r2+eps2) in marking order */
    i64 *leafcell;       /* opened leaf cells */
    i64 ntree, nleaf;
} twalk;

/* MarkCell (unbind.cxx:24-44) */
static void mark_cell(twalk *w, i64 nid, const double *xpos)
{
    const tnode *nd = &w->t->nodes[nid];
    double r2 = 0;
    for (int k = 0; k < 3; k++) r2 += (nd->cm[k] - xpos[k]) * (nd->cm[k] - xpos[k]);
    if (r2 < nd->r2max) {
        if (nd->left >= 0) {
            mark_cell(w, nd->left, xpos);
            mark_cell(w, nd->right, xpos);
        } else w->leafcell[w->nleaf++] = nid;
    } else {
        w->r2val[w->ntree++] = nd->mtot / sqrt(r2 + w->eps2);
    }
}

/* The arrangement BuildNodes leaves behind: perm_out[i] = input index of the particle at tree position i; node_out
 * (may be NULL, 2 entries per node, capacity node_cap nodes) receives [start, end) of every node in the pre-order of
 * GetNodeList.  Returns the number of nodes.  (Test hook for the device build's tie handling.) */
long long vro_tree_order(long long n, const double *pos, int bucket, long long *perm_out, long long *node_out, long long node_cap)
{
    if (n <= 0) return 0;
    tbuild t;
    t.pos = pos;
    t.bucket = bucket;
    t.perm = perm_out;
    for (i64 i = 0; i < n; i++) t.perm[i] = i;
    t.cap = 2 * n / (bucket > 0 ? bucket : 1) + 64;
    t.nodes = (tnode *)malloc(sizeof(tnode) * (size_t)t.cap);
    t.nnodes = 0;
    build_nodes(&t, 0, n);
    if (node_out)
        for (i64 j = 0; j < t.nnodes && j < node_cap; j++) { node_out[2 * j] = t.nodes[j].start; node_out[2 * j + 1] = t.nodes[j].end; }
    i64 nn = t.nnodes;
    free(t.nodes);
    return nn;
}

/* ---------------------------------------------------------------------------------------------------------------------
 * Row f4, oracle side: GetVelocityDensity (localfield.cxx:18) as SearchSubSub calls it for one
 * structure (tree == NULL): a physical kd-tree with bucket Options::Bsize and the Epanechnikov kernel table of 1000
 * entries (localfield.cxx:765, KDTree.cxx:592-631), then
 *   approx = 1 (GetVelocityDensityApproximative, localfield.cxx:726-897, the default): per LEAF of that tree the centre
 *     of its typed particles, the Nsearch nearest particles of that centre, and for every typed particle of the leaf the
 *     Nvel nearest of those in velocity (itself excluded);
 *   approx = 0 (GetVelocityDensityExact, localfield.cxx:485-586): the Nsearch nearest particles of every typed particle
 *     (itself excluded) and the Nvel nearest of those in velocity;
 * and KDTree::CalcSmoothLocalValue (KDCalcSmoothQuantities.cxx:1704-1719) on the velocity distances, largest first.
 * Neighbour sets are taken as the Nsearch smallest squared distances (the reference's tree search returns the same set
 * unless the Nsearch-th and (Nsearch+1)-th distances are exactly equal), periodic domains through the reference's own
 * sequence of reflected searches (nearest_pos).
 * ------------------------------------------------------------------------------------------------------------------- */
typedef struct { double d2; i64 id; } nnrec;
static int cmp_nnrec(const void *a, const void *b)
{
    const nnrec *x = (const nnrec *)a, *y = (const nnrec *)b;
    if (x->d2 < y->d2) return -1;
    if (x->d2 > y->d2) return 1;
    return (x->id > y->id) - (x->id < y->id);
}
static int cmp_dbl_desc(const void *a, const void *b)
{
    double x = *(const double *)a, y = *(const double *)b;
    return (x < y) - (x > y);
}
/* the K smallest records (by distance, then id) to the front of a[0..m), in order */
static void smallest_first(nnrec *a, i64 m, i64 K)
{
    if (K >= m) { qsort(a, (size_t)m, sizeof(nnrec), cmp_nnrec); return; }
    i64 lo = 0, hi = m - 1;
    while (lo < hi) { /* quickselect for position K - 1 */
        nnrec pv = a[lo + (hi - lo) / 2];
        i64 i = lo, j = hi;
        while (i <= j) {
            while (cmp_nnrec(&a[i], &pv) < 0) i++;
            while (cmp_nnrec(&a[j], &pv) > 0) j--;
            if (i <= j) { nnrec w = a[i]; a[i] = a[j]; a[j] = w; i++; j--; }
        }
        if (K - 1 <= j) hi = j;
        else if (K - 1 >= i) lo = i;
        else break;
    }
    qsort(a, (size_t)K, sizeof(nnrec), cmp_nnrec);
}
/* offer the squared distances of all particles to the point q to the list of the K smallest (work[0..K) sorted,
 * work[K..K+n) scratch) -- what one tree search does to the reference's priority queue */
static void offer_all(i64 n, const double *tp, const double *q, i64 K, nnrec *work)
{
    for (i64 i = 0; i < n; i++) {
        double d2 = 0;
        for (int k = 0; k < 3; k++) d2 += (q[k] - tp[3 * i + k]) * (q[k] - tp[3 * i + k]);
        work[K + i].d2 = d2;
        work[K + i].id = i;
    }
    smallest_first(work, K + n, K);
}
/* the nsearch nearest particles (tree order ids, -1 = none) of position c.
 * Open domain, target < 0: KDTree::FindNearestPos (KDFindNearest.cxx:559-571).  Open domain, target >= 0: KDTree::
 * FindNearest(target, ...) (:289-340), which leaves out the target and every particle at distance exactly zero
 * (KDLeafNode.cxx:19-21).
 * Periodic domain: SplitNode::FindNearestPosPeriodic (KDSplitNode.cxx:1008-1040) searches from the point itself and then
 * from up to seven reflected points, each search offering every particle to the same queue (so a particle can enter
 * twice, under two images); a reflection is searched only if the current largest queued distance reaches past the
 * faces it crosses -- for the one-face reflections sqrt(r2max) > distance, for the two- and three-face ones the test
 * compares the SQUARED radius with the distance (r2max > sqrt(sum of squared face distances)), reproduced here.  With a
 * target the queue holds one more entry and the smallest is dropped at the end (KDFindNearest.cxx:292, 356-357). */
static void nearest_pos(i64 n, const double *tp, const double *c, double period, i64 nsearch, nnrec *work, i64 *ids, i64 target)
{
    if (!(period > 0)) {
        i64 m = 0;
        for (i64 i = 0; i < n; i++) {
            if (i == target) continue;
            double d2 = 0;
            for (int k = 0; k < 3; k++) d2 += (c[k] - tp[3 * i + k]) * (c[k] - tp[3 * i + k]);
            if (target >= 0 && !(d2 > 0)) continue;
            work[m].d2 = d2;
            work[m].id = i;
            m++;
        }
        smallest_first(work, m, nsearch);
        for (i64 k = 0; k < nsearch; k++) ids[k] = k < m ? work[k].id : -1;
        return;
    }
    const i64 K = nsearch + (target >= 0 ? 1 : 0);
    for (i64 k = 0; k < K; k++) { work[k].d2 = 1e32; work[k].id = -1; } /* MAXVALUE, Precision.h:39 */
    offer_all(n, tp, c, K, work);
    double xp[3], sval;
    for (int k = 0; k < 3; k++) {                                  /* PeriodicReflection1D, DistFunc.h:326-332 */
        for (int j = 0; j < 3; j++) xp[j] = c[j];
        if (c[k] < period / 2.0) { xp[k] = c[k] + period; sval = c[k]; }
        else { xp[k] = c[k] - period; sval = -xp[k]; }
        if (sqrt(work[K - 1].d2) > sval) offer_all(n, tp, xp, K, work);
    }
    static const int pairs[3][2] = {{0, 1}, {0, 2}, {1, 2}};
    for (int q = 0; q <= 3; q++) {                                 /* PeriodicReflection2D x3, then ND (:338-355) */
        for (int j = 0; j < 3; j++) xp[j] = c[j];
        sval = 0;
        for (int a = 0; a < (q < 3 ? 2 : 3); a++) {
            const int k = q < 3 ? pairs[q][a] : a;
            if (c[k] < period / 2.0) { xp[k] = c[k] + period; sval += c[k] * c[k]; }
            else { xp[k] = c[k] - period; sval += xp[k] * xp[k]; }
        }
        sval = sqrt(sval);
        if (work[K - 1].d2 > sval) offer_all(n, tp, xp, K, work);
    }
    for (i64 k = 0; k < nsearch; k++) ids[k] = work[k + (K - nsearch)].id;
}
/* CalcSmoothLocalValue on nvel squared distances sorted largest first, unit weights, ND = 3 */
static double smooth_local_value(const double *v2desc, i64 nvel, const double *kernel, int kernres)
{
    double hi = 0.5 * sqrt(v2desc[0]);
    double norm = 1.0 / pow(hi, 3.0);
    double delta = 2.0 / (double)(kernres - 1);
    double value = 0;
    for (i64 j = 0; j < nvel; j++) {
        double rij = sqrt(v2desc[j]);
        double r = rij / hi;
        i64 i = (i64)(rij / hi * 0.5 * (kernres - 1));
        double w = i < kernres - 1 ? kernel[i] + (kernel[i + 1] - kernel[i]) * (r - delta * i) / delta : kernel[i];
        value += w * norm * 1.0;
    }
    return value;
}
long long vro_velocity_density(long long n, const double *pos, const double *vel, const int *type, double period, int bsize,
                               int nsearch, int nvel, int approx, double *den_out)
{
    for (i64 i = 0; i < n; i++) den_out[i] = 0.;
    if (n <= 0) return 0;
    if (nsearch > n - 1 || nvel > nsearch - 1) return -1;
    /* kernel table: KDTree::KernelConstruction, ND = 3, KEPAN, kernres = 1000 */
    const int kernres = 1000;
    double *kernel = (double *)malloc(sizeof(double) * kernres);
    {
        const int ND = 3;
        double kernnorm = 1.0 / ND * pow(M_1_PI, ND / 2.) * tgamma(ND / 2. + 1.);
        kernnorm *= ND * (ND + 2.) * pow(0.5, ND + 1.);
        double delta = 2.0 / (double)(kernres - 1);
        for (int i = 0; i < kernres; i++) {
            double rh = i * delta / 1.0 * 0.5;
            kernel[i] = kernnorm * (rh <= 1.0 ? (1.0 - rh * rh) : 0.);
        }
    }
    tbuild t;
    t.pos = pos;
    t.bucket = bsize;
    t.perm = (i64 *)malloc(sizeof(i64) * (size_t)n);
    for (i64 i = 0; i < n; i++) t.perm[i] = i;
    t.cap = 2 * n / (bsize > 0 ? bsize : 1) + 64;
    t.nodes = (tnode *)malloc(sizeof(tnode) * (size_t)t.cap);
    t.nnodes = 0;
    build_nodes(&t, 0, n);
    double *tp = (double *)malloc(sizeof(double) * 3 * (size_t)n), *tv = (double *)malloc(sizeof(double) * 3 * (size_t)n);
    int *tt = (int *)malloc(sizeof(int) * (size_t)n);
    for (i64 i = 0; i < n; i++) {
        for (int k = 0; k < 3; k++) { tp[3 * i + k] = pos[3 * t.perm[i] + k]; tv[3 * i + k] = vel[3 * t.perm[i] + k]; }
        tt[i] = type ? type[t.perm[i]] : 1;
    }
    nnrec *work = (nnrec *)malloc(sizeof(nnrec) * (size_t)(n + nsearch + 2));
    i64 *ids = (i64 *)malloc(sizeof(i64) * (size_t)nsearch);
    double *v2 = (double *)malloc(sizeof(double) * (size_t)nsearch);
    i64 nleaf = 0;
    for (i64 nd = 0; nd < t.nnodes; nd++) {
        if (t.nodes[nd].left >= 0) continue;
        nleaf++;
        const i64 s = t.nodes[nd].start, e = t.nodes[nd].end;
        if (approx) {
            i64 num = 0;
            double cm[3] = {0, 0, 0};
            for (i64 j = s; j < e; j++) if (tt[j] != 0) { num++; for (int k = 0; k < 3; k++) cm[k] += tp[3 * j + k]; }
            if (num == 0) continue;
            for (int k = 0; k < 3; k++) cm[k] /= (float)num;
            nearest_pos(n, tp, cm, period, nsearch, work, ids, -1);
        }
        for (i64 j = s; j < e; j++) {
            if (tt[j] == 0) continue;
            if (!approx) nearest_pos(n, tp, tp + 3 * j, period, nsearch, work, ids, j);
            i64 m = 0;
            for (i64 k = 0; k < nsearch; k++) {
                i64 id = ids[k];
                if (id < 0 || (approx && id == j)) continue;
                double d = 0;
                for (int c = 0; c < 3; c++) d += (tv[3 * j + c] - tv[3 * id + c]) * (tv[3 * j + c] - tv[3 * id + c]);
                v2[m++] = d;
            }
            /* the priority queue keeps the nvel smallest; CalcSmoothLocalValue pops them largest first */
            qsort(v2, (size_t)m, sizeof(double), cmp_dbl_desc);
            if (m < nvel) continue; /* fewer usable neighbours than Nvel (coincident particles): left at zero */
            den_out[t.perm[j]] = smooth_local_value(v2 + (m - nvel), nvel, kernel, kernres);
        }
    }
    free(v2); free(ids); free(work); free(tt); free(tv); free(tp); free(t.nodes); free(t.perm); free(kernel);
    return nleaf;
}

void vro_potential_tree(long long n, const double *pos, const double *mass, double G, double eps, double theta,
                        int bucket, double *phi_out, long long *interactions_out)
{
    if (n <= 0) return;
    tbuild t;
    t.pos = pos;
    t.bucket = bucket;
    t.perm = (i64 *)malloc(sizeof(i64) * (size_t)n);
    for (i64 i = 0; i < n; i++) t.perm[i] = i;
    t.cap = 2 * n / (bucket > 0 ? bucket : 1) + 64;
    t.nodes = (tnode *)malloc(sizeof(tnode) * (size_t)t.cap);
    t.nnodes = 0;
    build_nodes(&t, 0, n);
    double *tp = (double *)malloc(sizeof(double) * 3 * (size_t)n);
    double *tm = (double *)malloc(sizeof(double) * (size_t)n);
    for (i64 i = 0; i < n; i++) {
        for (int k = 0; k < 3; k++) tp[3 * i + k] = pos[3 * t.perm[i] + k];
        tm[i] = mass[t.perm[i]];
    }
    /* unbind.cxx:1667-1685 */
    for (i64 j = 0; j < t.nnodes; j++) {
        tnode *nd = &t.nodes[j];
        nd->cm[0] = nd->cm[1] = nd->cm[2] = 0.;
        nd->mtot = 0;
        for (i64 k = nd->start; k < nd->end; k++) {
            for (int d = 0; d < 3; d++) nd->cm[d] += tp[3 * k + d] * tm[k];
            nd->mtot += tm[k];
        }
        for (int d = 0; d < 3; d++) nd->cm[d] /= nd->mtot;
        double xdiff = 0;
        for (i64 k = nd->start; k < nd->end; k++) {
            /* Coordinate::Length(): sqrt(x*x+y*y+z*z) */
            double dx = nd->cm[0] - tp[3 * k], dy = nd->cm[1] - tp[3 * k + 1], dz = nd->cm[2] - tp[3 * k + 2];
            double l = sqrt(dx * dx + dy * dy + dz * dz);
            if (k == nd->start || xdiff < l) xdiff = l;
        }
        nd->r2max = 4.0 / 3.0 * xdiff * xdiff / (theta * theta);
    }
    twalk w;
    w.t = &t; w.tp = tp; w.tm = tm; w.eps2 = eps * eps;
    w.r2val = (double *)malloc(sizeof(double) * (size_t)t.nnodes);
    w.leafcell = (i64 *)malloc(sizeof(i64) * (size_t)t.nnodes);
    i64 inter = 0;
    /* unbind.cxx:1698-1724 */
    for (i64 j = 0; j < n; j++) {
        double phi = 0.;
        w.ntree = w.nleaf = 0;
        mark_cell(&w, 0, &tp[3 * j]);
        for (i64 k = 0; k < w.ntree; k++) phi = phi - tm[j] * w.r2val[k];
        inter += w.ntree;
        for (i64 k = 0; k < w.nleaf; k++) {
            const tnode *nd = &t.nodes[w.leafcell[k]];
            for (i64 l = nd->start; l < nd->end; l++) {
                if (j != l) {
                    double r2 = 0.;
                    for (int d = 0; d < 3; d++) r2 += pow(tp[3 * j + d] - tp[3 * l + d], 2.0);
                    r2 += w.eps2;
                    r2 = 1.0 / sqrt(r2);
                    phi = phi - (tm[j] * tm[l]) * r2;
                    inter++;
                }
            }
        }
        phi_out[t.perm[j]] = phi * G;
    }
    if (interactions_out) *interactions_out = inter;
    free(w.leafcell); free(w.r2val); free(tm); free(tp); free(t.nodes); free(t.perm);
}

/* ------------------------------------------------------------------------------------------------ */
/* unbinding                                                                                        */
/* ------------------------------------------------------------------------------------------------ */
typedef struct {
    i64 idx;       /* index into the caller's arrays (gPart[i][j].PID, unbind.cxx:231) */
    double E;      /* energy stored in rho (unbind.cxx:174) */
    double phi;
} upart;

static int cmp_upart(const void *a, const void *b)
{
    const upart *x = (const upart *)a, *y = (const upart *)b;
    if (x->E < y->E) return -1;
    if (x->E > y->E) return 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx);
}

typedef struct {
    const double *vel, *mass;
    const vro_unbind_opts *o;
    double marginal_tol;
    i64 n_marginal;
    unsigned char *marg; /* per particle marginal flag at last evaluation */
} uctx;

/* GetBoundFractionAndMaxE (unbind.cxx:146-185) */
static void bound_fraction(uctx *c, i64 ning, upart *gp, const double *cmvel, double *Efrac, double *maxE,
                           i64 *nunbound, int sortflag)
{
    double ef = 0;
    i64 nu = 0;
    double mx = -1e32; /* -MAXVALUE (Precision.h:40) */
    for (i64 j = 0; j < ning; j++) {
        i64 i = gp[j].idx;
        double m = c->mass[i];
        double v2 = 0.0;
        for (int k = 0; k < 3; k++) v2 += pow(c->vel[3 * i + k] - cmvel[k], 2.0);
        double Ti = 0.5 * m * v2;
        gp[j].E = c->o->Eratio * Ti + gp[j].phi;
        if (gp[j].E > mx) mx = gp[j].E;
        ef += (Ti + gp[j].phi < 0);
        if (gp[j].E > 0) nu++;
        c->marg[i] = fabs(gp[j].E) <= c->marginal_tol * fabs(gp[j].phi);
    }
    ef /= (double)ning;
    if (nu < c->o->maxunboundfracforiterativeunbind * ning && sortflag) qsort(gp, (size_t)ning, sizeof(upart), cmp_upart);
    *Efrac = ef; *maxE = mx; *nunbound = nu;
}

/* CheckGroupForBoundness (unbind.cxx:49-60) */
static int check_bound(const vro_unbind_opts *o, double Efrac, double maxE, i64 ning)
{
    if (o->unbindtype == 0) return ((Efrac < o->minEfrac) || (maxE > 0)) && (ning >= o->min_size);
    return (maxE > 0) && (ning >= o->min_size);
}

/* FillUnboundArrays (unbind.cxx:62-98); returns nEplus, may clear *unbindcheck */
static i64 fill_unbound(const vro_unbind_opts *o, i64 maxunbindsize, i64 ning, upart *gp, double Efrac,
                        i64 *nEplusid, int *unbindcheck)
{
    i64 nEplus = 0;
    if (!*unbindcheck) return 0;
    if (o->unbindtype == 1) {
        for (i64 j = 0; j < maxunbindsize; j++) {
            if (ning - 1 - j < 0) break;
            if (gp[ning - 1 - j].E > 0) nEplusid[nEplus++] = ning - 1 - j;
            else break;
        }
    } else {
        i64 nEfrac = 0;
        if (Efrac < o->minEfrac) nEfrac = (i64)((o->minEfrac - Efrac) * ning);
        for (i64 j = 0; j < maxunbindsize; j++) {
            if (ning - 1 - j < 0) break;
            if (gp[ning - 1 - j].E > 0 || nEplus < nEfrac) nEplusid[nEplus++] = ning - 1 - j;
            else break;
        }
    }
    if (nEplus < o->maxallowedunboundfrac * ning) { nEplus = 0; *unbindcheck = 0; }
    return nEplus;
}

/* one pass of the unbinding loop body shared by the large (unbind.cxx:552-639) and small (:641-720) loops.
 * `large` selects the termination rule (quirk Q9). */
static void unbind_group(uctx *c, i64 *ning_io, upart *gp, double *cmvel, double *gmass, const double *pos,
                         i64 *pfof, int large, i64 *oldnumingroup, int *iunbindflag)
{
    const vro_unbind_opts *o = c->o;
    i64 ning = *ning_io;
    double Efrac, maxE;
    i64 nunbound;
    if (large || o->sortflag_mode == 0) *oldnumingroup = ning; /* unbind.cxx:555 (small loop: never set, Q10) */
    bound_fraction(c, ning, gp, cmvel, &Efrac, &maxE, &nunbound, 1);
    if (nunbound >= o->maxunboundfracforiterativeunbind * ning) {
        for (i64 j = 0; j < ning; j++) pfof[gp[j].idx] = 0;
        *ning_io = 0;
        (*iunbindflag)++;
        return;
    }
    i64 maxunbindsize = (i64)(o->maxunbindfrac * nunbound + 1);
    i64 *nEplusid = (i64 *)malloc(sizeof(i64) * (size_t)(ning + 1));
    int unbindcheck = check_bound(o, Efrac, maxE, ning);
    i64 nEplus = fill_unbound(o, maxunbindsize, ning, gp, Efrac, nEplusid, &unbindcheck);
    double eps2 = o->eps * o->eps;
    while (unbindcheck) {
        (*iunbindflag)++;
        /* UpdateCMForUnboundParticles (unbind.cxx:111-127): only the whole-group frame follows the removals */
        if (o->fracpotref == 1.0) {
            double temp = 1.0 / (*gmass), temp2 = 0.;
            for (i64 j = 0; j < nEplus; j++) {
                i64 i = gp[nEplusid[j]].idx;
                for (int k = 0; k < 3; k++) cmvel[k] -= c->vel[3 * i + k] * c->mass[i] * temp;
                temp2 += c->mass[i];
            }
            temp = (*gmass) / ((*gmass) - temp2);
            for (int k = 0; k < 3; k++) cmvel[k] *= temp;
            *gmass -= temp2;
        }
        /* unbind.cxx:587-605 / 674-688: take off the pair terms of every removed particle.  The large-group loop does
         * that only for nEplus < 2 ln n (unbind.cxx:583); otherwise it calls Potential() on the member list BEFORE the
         * removal (unbind.cxx:585 precedes RemoveUnboundParticles at :607), i.e. it re-evaluates the potentials the
         * members already carry, so nothing changes here. */
        if (o->bgpot == 0 && (!large || (double)nEplus < 2.0 * log((double)ning))) {
            for (i64 k = 0; k < nEplus; k++) {
                i64 ik = gp[nEplusid[k]].idx;
                for (i64 j = 0; j < ning; j++) {
                    if (j != nEplusid[k]) {
                        i64 ij = gp[j].idx;
                        double r2 = 0.;
                        for (int d = 0; d < 3; d++) r2 += pow(pos[3 * ik + d] - pos[3 * ij + d], 2.0);
                        r2 += eps2;
                        r2 = 1.0 / sqrt(r2);
                        gp[j].phi = gp[j].phi + o->G * (c->mass[ik] * c->mass[ij]) * r2;
                    }
                }
            }
        }
        /* RemoveUnboundParticles (unbind.cxx:101-108) */
        for (i64 j = 0; j < nEplus; j++) pfof[gp[nEplusid[j]].idx] = 0;
        ning -= nEplus;
        int stop = large ? (nEplus > o->maxallowedunboundfrac * ning) : (nEplus < o->maxallowedunboundfrac * ning);
        if (stop) { unbindcheck = 0; continue; }
        int sortflag = 0;
        if ((*oldnumingroup - ning) > o->maxallowedunboundfrac * (*oldnumingroup)) { *oldnumingroup = ning; sortflag = 1; }
        bound_fraction(c, ning, gp, cmvel, &Efrac, &maxE, &nunbound, sortflag);
        maxunbindsize = (i64)(o->maxunbindfrac * nunbound + 1);
        unbindcheck = check_bound(o, Efrac, maxE, ning);
        nEplus = fill_unbound(o, maxunbindsize, ning, gp, Efrac, nEplusid, &unbindcheck);
    }
    /* RemoveGroup (unbind.cxx:129-140) */
    if (ning < o->min_size) {
        for (i64 j = 0; j < ning; j++) pfof[gp[j].idx] = 0;
        ning = 0;
        (*iunbindflag)++;
    }
    free(nEplusid);
    *ning_io = ning;
}

typedef struct { double r; i64 j; } radrec;
static int cmp_radrec(const void *a, const void *b)
{
    const radrec *x = (const radrec *)a, *y = (const radrec *)b;
    if (x->r != y->r) return x->r < y->r ? -1 : 1;
    return x->j < y->j ? -1 : (x->j > y->j ? 1 : 0);
}

long long vro_unbind(long long n, const double *pos, const double *vel, const double *mass, long long *pfof,
                     long long ngroups, const vro_unbind_opts *o, const long long *pid, double marginal_tol,
                     long long *n_marginal_out, double *phi_out, int *changed_out)
{
    if (changed_out) *changed_out = 0;
    if (n_marginal_out) *n_marginal_out = 0;
    if (ngroups <= 0) return 0;
    /* BuildNumInGroup / BuildPGList (buildandsortarrays.cxx:11-43): members in index order */
    i64 *num = (i64 *)calloc((size_t)ngroups + 2, sizeof(i64));
    i64 *off = (i64 *)calloc((size_t)ngroups + 2, sizeof(i64));
    for (i64 i = 0; i < n; i++) if (pfof[i] > 0) num[pfof[i]]++;
    for (i64 g = 2; g <= ngroups + 1; g++) off[g] = off[g - 1] + num[g - 1];
    i64 ntot = off[ngroups + 1];
    upart *gp = (upart *)malloc(sizeof(upart) * (size_t)(ntot + 1));
    {
        i64 *fill = (i64 *)calloc((size_t)ngroups + 2, sizeof(i64));
        for (i64 i = 0; i < n; i++) if (pfof[i] > 0) { i64 g = pfof[i]; gp[off[g] + fill[g]++].idx = i; }
        free(fill);
    }
    uctx c;
    c.vel = vel; c.mass = mass; c.o = o; c.marginal_tol = marginal_tol; c.n_marginal = 0;
    c.marg = (unsigned char *)calloc((size_t)n + 1, 1);
    double *cmvel = (double *)calloc(3 * ((size_t)ngroups + 2), sizeof(double));
    double *gmass = (double *)calloc((size_t)ngroups + 2, sizeof(double));
    /* potentials: unbind.cxx:398-434 */
    i64 nmax = 0;
    for (i64 g = 1; g <= ngroups; g++) if (num[g] > nmax) nmax = num[g];
    double *gx = (double *)malloc(sizeof(double) * 3 * (size_t)(nmax + 1));
    double *gm = (double *)malloc(sizeof(double) * (size_t)(nmax + 1));
    double *gphi = (double *)malloc(sizeof(double) * (size_t)(nmax + 1));
    for (i64 g = 1; g <= ngroups; g++) {
        i64 ng = num[g];
        upart *p = gp + off[g];
        for (i64 j = 0; j < ng; j++) {
            for (int k = 0; k < 3; k++) gx[3 * j + k] = pos[3 * p[j].idx + k];
            gm[j] = mass[p[j].idx];
        }
        if (o->pp_threshold < 0 || ng <= o->pp_threshold) vro_potential_pp(ng, gx, gm, o->G, o->eps, gphi);
        else vro_potential_tree(ng, gx, gm, o->G, o->eps, o->theta, o->bucket, gphi, NULL);
        for (i64 j = 0; j < ng; j++) { p[j].phi = gphi[j]; p[j].E = 0; if (phi_out) phi_out[p[j].idx] = gphi[j]; }
        if (o->fracpotref == 1.0) {
            /* unbind.cxx:447-455 */
            for (i64 j = 0; j < ng; j++) {
                gmass[g] += mass[p[j].idx];
                for (int k = 0; k < 3; k++) cmvel[3 * g + k] += vel[3 * p[j].idx + k] * mass[p[j].idx];
            }
            double inv = 1.0 / gmass[g];
            for (int k = 0; k < 3; k++) cmvel[3 * g + k] *= inv;
        } else {
            /* unbind.cxx:460-544: velocity frame of the innermost npot members, by radius about the centre of mass (CMVELREF) or
             * about the member with the lowest potential (POTREF, first strict minimum).  gsl_heapsort leaves equal radii in an
             * order of its own; here they keep member order. */
            double c0[3] = {0., 0., 0.};
            if (o->cmvelreftype == 0) {
                for (i64 j = 0; j < ng; j++) {
                    gmass[g] += mass[p[j].idx];
                    for (int k = 0; k < 3; k++) c0[k] += pos[3 * p[j].idx + k] * mass[p[j].idx];
                }
                const double inv = 1.0 / gmass[g];
                for (int k = 0; k < 3; k++) c0[k] *= inv;
            } else {
                i64 imin = 0;
                for (i64 j = 1; j < ng; j++) if (p[j].phi < p[imin].phi) imin = j;
                for (int k = 0; k < 3; k++) c0[k] = pos[3 * p[imin].idx + k];
            }
            radrec *rr = (radrec *)malloc(sizeof(radrec) * (size_t)(ng + 1));
            for (i64 j = 0; j < ng; j++) {
                const double x = pos[3 * p[j].idx] - c0[0], y = pos[3 * p[j].idx + 1] - c0[1], z = pos[3 * p[j].idx + 2] - c0[2];
                rr[j].r = sqrt(x * x + y * y + z * z); /* Particle::Radius() */
                rr[j].j = j;
            }
            qsort(rr, (size_t)ng, sizeof(radrec), cmp_radrec);
            i64 npot = (i64)(o->fracpotref * (double)ng);
            if (npot < o->npotref) npot = o->npotref;
            if (npot > ng) npot = ng;
            double menc = 0.;
            for (i64 q = 0; q < npot; q++) {
                const i64 i = p[rr[q].j].idx;
                for (int k = 0; k < 3; k++) cmvel[3 * g + k] += vel[3 * i + k] * mass[i];
                menc += mass[i];
            }
            for (int k = 0; k < 3; k++) cmvel[3 * g + k] /= menc;
            free(rr);
        }
    }
    free(gphi); free(gm); free(gx);

    int iunbindflag = 0;
    i64 oldnumingroup = o->sortflag_mode == 1 ? o->oldn_init : 0;
    /* unbind.cxx:552: large groups first, serial */
    for (i64 g = 1; g <= ngroups; g++)
        if (num[g] >= o->large_threshold)
            unbind_group(&c, &num[g], gp + off[g], &cmvel[3 * g], &gmass[g], pos, pfof, 1, &oldnumingroup, &iunbindflag);
    /* unbind.cxx:647: then every group that is now <= ompunbindnum and non-empty */
    for (i64 g = 1; g <= ngroups; g++)
        if (num[g] <= o->large_threshold && num[g] > 0)
            unbind_group(&c, &num[g], gp + off[g], &cmvel[3 * g], &gmass[g], pos, pfof, 0, &oldnumingroup, &iunbindflag);

    i64 ng = ngroups;
    for (i64 g = 1; g <= ngroups; g++) if (num[g] == 0) ng--;
    if (n_marginal_out) {
        i64 nm = 0;
        for (i64 i = 0; i < n; i++) nm += c.marg[i];
        *n_marginal_out = nm;
    }
    if (changed_out) *changed_out = iunbindflag ? 1 : 0;
    /* ReorderGroupIDs (unbind.cxx:722, buildandsortarrays.cxx:170-181) -> canonical order */
    if (iunbindflag && ng > 0) {
        grec *gr = (grec *)malloc(sizeof(grec) * (size_t)(ngroups + 1));
        for (i64 k = 0; k <= ngroups; k++) { gr[k].size = 0; gr[k].minpid = 0x7fffffffffffffffLL; gr[k].root = k; }
        for (i64 i = 0; i < n; i++) if (pfof[i] > 0) {
            grec *r = &gr[pfof[i]];
            r->size++;
            i64 p = pid ? pid[i] : i;
            if (p < r->minpid) r->minpid = p;
        }
        qsort(gr + 1, (size_t)ngroups, sizeof(grec), cmp_grec);
        i64 *newid = (i64 *)calloc((size_t)ngroups + 1, sizeof(i64));
        for (i64 k = 1; k <= ngroups; k++) newid[gr[k].root] = gr[k].size > 0 ? k : 0;
        for (i64 i = 0; i < n; i++) if (pfof[i] > 0) pfof[i] = newid[pfof[i]];
        free(newid);
        free(gr);
    } else if (ng == 0) {
        for (i64 i = 0; i < n; i++) pfof[i] = 0;
    }
    free(gmass); free(cmvel); free(c.marg); free(gp); free(off); free(num);
    return ng;
}


/* ------------------------------------------------------------------------------------------------------
 * GetBindingEnergy (substructureproperties.cxx:3884-4255), default reference frame.  See vr_oracle.h.
 * ------------------------------------------------------------------------------------------------------ */
void vro_binding_energy(long long n, const double *pos, const double *vel, const double *mass, long long ngroups,
                        const long long *off, const double *gcmvel, double G, double eps, double theta, int bucket,
                        long long large_threshold, double *phi, double *E, double *T_out, double *Pot_out, double *Efrac_out,
                        long long *iminpot_out, long long *ibound_out)
{
    const double eps2 = eps * eps;
    for (i64 j = 0; j < n; j++) { phi[j] = 0.0; E[j] = 0.0; }
    for (i64 g = 1; g <= ngroups; g++) {
        const i64 o = off[g], ng = off[g + 1] - off[g];
        double Pot = 0.0, Tsum = 0.0, Efrac = 0.0;
        if (ng <= 0) { T_out[g] = Pot_out[g] = Efrac_out[g] = 0.0; iminpot_out[g] = ibound_out[g] = 0; continue; }
        const int large = ng >= large_threshold;
        if (!large) {
            /* :3922-3940 pair loop, potentials accumulate in place, pdata.Pot sums the pairs */
            for (i64 j = 0; j < ng; j++)
                for (i64 k = j + 1; k < ng; k++) {
                    double r2 = 0.;
                    for (int d = 0; d < 3; d++) r2 += pow(pos[3 * (o + j) + d] - pos[3 * (o + k) + d], 2.0);
                    r2 += eps2;
                    r2 = 1.0 / sqrt(r2);
                    double pot = -G * (mass[o + j] * mass[o + k]) * r2;
                    Pot += pot;
                    phi[o + j] = phi[o + j] + pot;
                    phi[o + k] = phi[o + k] + pot;
                }
        } else {
            /* :4060-4073 Potential() on the group's slice */
            vro_potential_tree(ng, pos + 3 * o, mass + o, G, eps, theta, bucket, phi + o, NULL);
        }
        /* minimum potential particle: first strict minimum (:4003-4016 / :4150-4163) */
        {
            double potmin = phi[o];
            i64 ipotmin = 0;
            for (i64 j = 0; j < ng; j++)
                if (phi[o + j] < potmin) { potmin = phi[o + j]; ipotmin = j; }
            iminpot_out[g] = ipotmin;
        }
        /* kinetic energies in the frame GetCM left in gcmvel (:4025-4052 small, :4166-4209 large) */
        for (i64 j = 0; j < ng; j++) {
            double v2 = 0.;
            for (int d = 0; d < 3; d++) v2 += pow(vel[3 * (o + j) + d] - gcmvel[3 * g + d], 2.0);
            double Ti = 0.5 * mass[o + j] * v2;
            if (large) {
                Pot += phi[o + j];
                E[o + j] = phi[o + j] + Ti;
            } else
                E[o + j] = Ti + phi[o + j];
            Tsum += Ti;
            if (E[o + j] < 0.0) Efrac += 1.0;
        }
        T_out[g] = Tsum;
        Pot_out[g] = Pot;
        Efrac_out[g] = Efrac / (double)ng;
        /* most bound particle: first strict minimum of E (:4216-4226) */
        {
            double Emost = E[o];
            i64 imost = 0;
            for (i64 j = 0; j < ng; j++)
                if (E[o + j] < Emost) { Emost = E[o + j]; imost = j; }
            ibound_out[g] = imost;
        }
    }
}
