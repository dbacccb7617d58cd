/* oracle/gsl_shim/gsl/gsl_shim_all.h
 *
 * TEST INFRASTRUCTURE ONLY.  Minimal stand-in for the handful of GSL entry points the
 * VELOCIraptor reference sources touch, so that the *unmodified* reference under
 * /root/reference can be compiled into oracle/_ref/ without a GSL install (none is
 * available in this image, and there is no network).
 *
 * What is real here:
 *   gsl_heapsort      - a genuine in-place heapsort over opaque elements (used on the hot
 *                       path by search.cxx:693 and unbind.cxx:482-537; with distinct keys any
 *                       correct sort yields the same permutation)
 *   gsl_sf_gamma      - tgamma (KDTree.cxx:606 kernel normalisation)
 *   gsl_sf_erf        - erf    (search.cxx:2964)
 *   gsl_integration_qags - composite Simpson (only CalcCosmicTime, off the hot path)
 *   gsl_root_fsolver_*   - bisection (only CalcConcentration, off the hot path)
 * What aborts if ever reached (none of it is on the FOF/unbinding path):
 *   gsl_rng_*, gsl_ran_gaussian, gsl_monte_vegas_*, gsl_cdf_chisq_*
 */
#ifndef VR_ORACLE_GSL_SHIM_ALL_H
#define VR_ORACLE_GSL_SHIM_ALL_H

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSL_SUCCESS 0
#define GSL_CONTINUE (-2)
#define GSL_FAILURE (-1)

#ifndef M_PI
#define M_PI 3.14159265358979323846264338328
#endif
#ifndef M_1_PI
#define M_1_PI 0.31830988618379067153776752675
#endif

/* ---- functions ---- */
typedef struct gsl_function_struct {
    double (*function)(double x, void *params);
    void *params;
} gsl_function;
#define GSL_FN_EVAL(F, x) (*((F)->function))(x, (F)->params)

typedef struct gsl_monte_function_struct {
    double (*f)(double *x_array, size_t dim, void *params);
    size_t dim;
    void *params;
} gsl_monte_function;

/* ---- heapsort ---- */
typedef int (*gsl_comparison_fn_t)(const void *, const void *);

static inline void vr_shim_swap_(char *a, char *b, size_t size)
{
    while (size--) { char t = *a; *a++ = *b; *b++ = t; }
}
static inline void vr_shim_sift_(char *base, size_t size, size_t n_last, size_t k, gsl_comparison_fn_t cmp)
{
    /* n_last is the index of the last live element */
    while (k <= n_last / 2) {
        size_t j = 2 * k;
        if (j < n_last && cmp(base + size * j, base + size * (j + 1)) < 0) j++;
        if (cmp(base + size * k, base + size * j) < 0) vr_shim_swap_(base + size * j, base + size * k, size);
        else break;
        k = j;
    }
}
static inline void gsl_heapsort(void *array, size_t count, size_t size, gsl_comparison_fn_t compare)
{
    if (count == 0) return;
    char *base = (char *)array;
    size_t n_last = count - 1;
    size_t k = n_last / 2;
    k++;
    do { k--; vr_shim_sift_(base, size, n_last, k, compare); } while (k > 0);
    while (n_last > 0) {
        vr_shim_swap_(base, base + n_last * size, size);
        n_last--;
        vr_shim_sift_(base, size, n_last, 0, compare);
    }
}

/* ---- special functions ---- */
static inline double gsl_sf_gamma(double x) { return tgamma(x); }
static inline double gsl_sf_erf(double x) { return erf(x); }

/* ---- integration (Simpson; off the hot path) ---- */
typedef struct { size_t limit; } gsl_integration_workspace;
static inline gsl_integration_workspace *gsl_integration_workspace_alloc(size_t n)
{
    gsl_integration_workspace *w = (gsl_integration_workspace *)malloc(sizeof(*w));
    w->limit = n;
    return w;
}
static inline void gsl_integration_workspace_free(gsl_integration_workspace *w) { free(w); }
static inline int gsl_integration_qags(const gsl_function *f, double a, double b, double epsabs, double epsrel,
                                       size_t limit, gsl_integration_workspace *w, double *result, double *abserr)
{
    (void)epsabs; (void)epsrel; (void)limit; (void)w;
    const int n = 4096; /* even */
    double h = (b - a) / n, s = GSL_FN_EVAL(f, a) + GSL_FN_EVAL(f, b);
    for (int i = 1; i < n; i++) s += GSL_FN_EVAL(f, a + i * h) * ((i & 1) ? 4.0 : 2.0);
    *result = s * h / 3.0;
    if (abserr) *abserr = 0.0;
    return GSL_SUCCESS;
}

/* ---- root finding (bisection; off the hot path) ---- */
typedef struct { const char *name; } gsl_root_fsolver_type;
typedef struct { gsl_function *F; double lo, hi, root; } gsl_root_fsolver;
static const gsl_root_fsolver_type vr_shim_brent_ = {"bisection-shim"};
static const gsl_root_fsolver_type *const gsl_root_fsolver_brent = &vr_shim_brent_;
static inline gsl_root_fsolver *gsl_root_fsolver_alloc(const gsl_root_fsolver_type *T)
{
    (void)T;
    return (gsl_root_fsolver *)calloc(1, sizeof(gsl_root_fsolver));
}
static inline int gsl_root_fsolver_set(gsl_root_fsolver *s, gsl_function *F, double lo, double hi)
{
    s->F = F; s->lo = lo; s->hi = hi; s->root = 0.5 * (lo + hi);
    return GSL_SUCCESS;
}
static inline int gsl_root_fsolver_iterate(gsl_root_fsolver *s)
{
    double flo = GSL_FN_EVAL(s->F, s->lo), mid = 0.5 * (s->lo + s->hi), fm = GSL_FN_EVAL(s->F, mid);
    if ((flo < 0) == (fm < 0)) s->lo = mid; else s->hi = mid;
    s->root = 0.5 * (s->lo + s->hi);
    return GSL_SUCCESS;
}
static inline double gsl_root_fsolver_root(const gsl_root_fsolver *s) { return s->root; }
static inline double gsl_root_fsolver_x_lower(const gsl_root_fsolver *s) { return s->lo; }
static inline double gsl_root_fsolver_x_upper(const gsl_root_fsolver *s) { return s->hi; }
static inline void gsl_root_fsolver_free(gsl_root_fsolver *s) { free(s); }
static inline int gsl_root_test_interval(double lo, double hi, double epsabs, double epsrel)
{
    double alo = fabs(lo), ahi = fabs(hi);
    double minabs = ((lo > 0 && hi > 0) || (lo < 0 && hi < 0)) ? (alo < ahi ? alo : ahi) : 0.0;
    return (fabs(hi - lo) < epsabs + epsrel * minabs) ? GSL_SUCCESS : GSL_CONTINUE;
}

/* ---- things that must never be reached on the path under test ---- */
static inline void vr_shim_die_(const char *what)
{
    fprintf(stderr, "gsl shim: %s is not available in the oracle build\n", what);
    abort();
}
typedef struct { int dummy; } gsl_rng_type;
typedef struct { int dummy; } gsl_rng;
static const gsl_rng_type vr_shim_rng_default_ = {0};
static const gsl_rng_type *const gsl_rng_default = &vr_shim_rng_default_;
static inline const gsl_rng_type *gsl_rng_env_setup(void) { return gsl_rng_default; }
static inline gsl_rng *gsl_rng_alloc(const gsl_rng_type *T) { (void)T; vr_shim_die_("gsl_rng_alloc"); return NULL; }
static inline void gsl_rng_set(const gsl_rng *r, unsigned long s) { (void)r; (void)s; vr_shim_die_("gsl_rng_set"); }
static inline void gsl_rng_free(gsl_rng *r) { (void)r; }
static inline double gsl_ran_gaussian(const gsl_rng *r, double sigma) { (void)r; (void)sigma; vr_shim_die_("gsl_ran_gaussian"); return 0; }

typedef struct { size_t dim; double chisq; } gsl_monte_vegas_state;
static inline gsl_monte_vegas_state *gsl_monte_vegas_alloc(size_t dim) { (void)dim; vr_shim_die_("gsl_monte_vegas_alloc"); return NULL; }
static inline int gsl_monte_vegas_integrate(gsl_monte_function *f, double *xl, double *xu, size_t dim, size_t calls,
                                            gsl_rng *r, gsl_monte_vegas_state *s, double *result, double *abserr)
{
    (void)f; (void)xl; (void)xu; (void)dim; (void)calls; (void)r; (void)s; (void)result; (void)abserr;
    vr_shim_die_("gsl_monte_vegas_integrate");
    return GSL_FAILURE;
}
static inline void gsl_monte_vegas_free(gsl_monte_vegas_state *s) { (void)s; }

static inline double gsl_cdf_chisq_Pinv(double P, double nu) { (void)P; (void)nu; vr_shim_die_("gsl_cdf_chisq_Pinv"); return 0; }
static inline double gsl_cdf_chisq_P(double x, double nu) { (void)x; (void)nu; vr_shim_die_("gsl_cdf_chisq_P"); return 0; }
static inline double gsl_cdf_chisq_Q(double x, double nu) { (void)x; (void)nu; vr_shim_die_("gsl_cdf_chisq_Q"); return 0; }

#ifdef __cplusplus
}
#endif
#endif
