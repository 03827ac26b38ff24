/*
 * bwk_oracle.c -- TEST INFRASTRUCTURE ONLY (see bwk_oracle.h).
 *
 * CPU restatement of atmosphere/biharmonic_wk_kernel.F90 (F90:<line> below).
 * Build for parity with  gcc -O2 -ffp-contract=off  so that no FMA contraction or
 * reassociation changes the reference's operation order.
 */
#include "bwk_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NP 4
#define ED BWKO_ELEM_DOUBLES
/* 1-based Fortran index helpers onto the flattened layouts */
#define DVV(d, i, l) (d)[((i)-1) + 4 * ((l)-1)]
#define S2(s, i, j) (s)[((i)-1) + 4 * ((j)-1)]
#define S3(s, i, j, c) (s)[((i)-1) + 4 * ((j)-1) + 16 * ((c)-1)]
#define DINV(e, i, j, c, a) (e)[((i)-1) + 4 * ((j)-1) + 16 * ((c)-1) + 32 * ((a)-1)]
#define SPMP(e, i, j) (e)[64 + ((i)-1) + 4 * ((j)-1)]
#define TVISC(e, i, j, c, a) (e)[80 + ((i)-1) + 4 * ((j)-1) + 16 * ((c)-1) + 32 * ((a)-1)]

/* F90:14: `rrearth = 0.00000016666666666666` is a default-real literal. */
double bwko_rrearth(void) { return (double)0.00000016666666666666f; }

/* ------------------------------------------------------------------ F90:77-91 */
void bwko_myrandom(int64_t n, double *a, int reset, int *state)
{
    int old = *state;
    if (reset) old = 11;
    for (int64_t i = 0; i < n; ++i) {
        old = (1301 * old + 97) % (1024 * 128);
        a[i] = old / (double)(1024 * 128);
    }
    *state = old;
}

/* ------------------------------------------------------------------ F90:48-58 */
void bwko_initialize_data(int nelemd, int nlev, int qsize, double *dvv, double *elem, double *qtens)
{
    int st = 11;
    bwko_myrandom(16, dvv, 1, &st);
    for (int ie = 0; ie < nelemd; ++ie) {
        bwko_myrandom(64, elem + (size_t)ED * ie, 0, &st);      /* Dinv       */
        bwko_myrandom(16, elem + (size_t)ED * ie + 64, 0, &st); /* spheremp   */
        bwko_myrandom(64, elem + (size_t)ED * ie + 80, 0, &st); /* tensorVisc */
    }
    bwko_myrandom((int64_t)16 * nlev * qsize * nelemd, qtens, 0, &st);
}

void bwko_initialize_subset(int64_t nelemd_total, int nlev, int qsize, int64_t first, int64_t count,
                            double *dvv, double *elem, double *qtens)
{
    enum { P = 131072 };
    double *table = (double *)malloc(sizeof(double) * P);
    int st = 11;
    bwko_myrandom(P, table, 1, &st); /* table[t] = draw number t+1 */
    for (int t = 0; t < 16; ++t) dvv[t] = table[t];
    for (int64_t e = 0; e < count; ++e) {
        int64_t d0 = 16 + (int64_t)ED * (first + e); /* 0-based draw index of Dinv(1,1,1,1) */
        for (int t = 0; t < ED; ++t) elem[ED * e + t] = table[(d0 + t) % P];
    }
    int64_t per_elem = (int64_t)16 * nlev * qsize;
    int64_t q0 = 16 + (int64_t)ED * nelemd_total + per_elem * first;
    for (int64_t t = 0; t < per_elem * count; ++t) qtens[t] = table[(q0 + t) % P];
    free(table);
}

/* ---------------------------------------------------------------- F90:109-134 */
static void gradient_sphere_cpu(const double *s, const double *dvv, const double *e, double *ds)
{
    const double rrearth = bwko_rrearth();
    double v1[16], v2[16];
    for (int j = 1; j <= NP; ++j) {
        for (int l = 1; l <= NP; ++l) {
            double dsdx00 = 0.0, dsdy00 = 0.0;
            for (int i = 1; i <= NP; ++i) {
                dsdx00 = dsdx00 + DVV(dvv, i, l) * S2(s, i, j);
                dsdy00 = dsdy00 + DVV(dvv, i, l) * S2(s, j, i);
            }
            S2(v1, l, j) = dsdx00 * rrearth;
            S2(v2, j, l) = dsdy00 * rrearth;
        }
    }
    for (int j = 1; j <= NP; ++j) {
        for (int i = 1; i <= NP; ++i) {
            S3(ds, i, j, 1) = DINV(e, i, j, 1, 1) * S2(v1, i, j) + DINV(e, i, j, 2, 1) * S2(v2, i, j);
            S3(ds, i, j, 2) = DINV(e, i, j, 1, 2) * S2(v1, i, j) + DINV(e, i, j, 2, 2) * S2(v2, i, j);
        }
    }
}

/* ---------------------------------------------------------------- F90:138-160 */
static void divergence_sphere_wk_cpu(const double *v, const double *dvv, const double *e, double *div)
{
    const double rrearth = bwko_rrearth();
    double vtemp[32];
    for (int j = 1; j <= NP; ++j) {
        for (int i = 1; i <= NP; ++i) {
            S3(vtemp, i, j, 1) = (DINV(e, i, j, 1, 1) * S3(v, i, j, 1) + DINV(e, i, j, 1, 2) * S3(v, i, j, 2));
            S3(vtemp, i, j, 2) = (DINV(e, i, j, 2, 1) * S3(v, i, j, 1) + DINV(e, i, j, 2, 2) * S3(v, i, j, 2));
        }
    }
    for (int n = 1; n <= NP; ++n) {
        for (int m = 1; m <= NP; ++m) {
            double d = 0.0;
            for (int j = 1; j <= NP; ++j) {
                d = d - (SPMP(e, j, n) * S3(vtemp, j, n, 1) * DVV(dvv, m, j) +
                         SPMP(e, m, j) * S3(vtemp, m, j, 2) * DVV(dvv, n, j)) * rrearth;
            }
            S2(div, m, n) = d;
        }
    }
}

/* ---------------------------------------------------------------- F90:164-182 */
static void laplace_sphere_wk_cpu(const double *s, const double *dvv, const double *e, double *laplace)
{
    double grads[32], oldgrads[32];
    gradient_sphere_cpu(s, dvv, e, grads);
    memcpy(oldgrads, grads, sizeof grads);
    for (int j = 1; j <= NP; ++j) {
        for (int i = 1; i <= NP; ++i) {
            S3(grads, i, j, 1) = S3(oldgrads, i, j, 1) * TVISC(e, i, j, 1, 1) + S3(oldgrads, i, j, 2) * TVISC(e, i, j, 1, 2);
            S3(grads, i, j, 2) = S3(oldgrads, i, j, 1) * TVISC(e, i, j, 2, 1) + S3(oldgrads, i, j, 2) * TVISC(e, i, j, 2, 2);
        }
    }
    divergence_sphere_wk_cpu(grads, dvv, e, laplace);
}

/* ---------------------------------------------------------------- F90:186-200 */
static void element_cpu(const double *e, double *qe, const double *dvv, int nlev, int qsize)
{
    for (int q = 0; q < qsize; ++q) {
        for (int k = 0; k < nlev; ++k) {
            double *slab = qe + (size_t)16 * (k + (size_t)nlev * q);
            double s[16], lap[16];
            memcpy(s, slab, sizeof s); /* the dummy argument copy-in of F90:196 */
            laplace_sphere_wk_cpu(s, dvv, e, lap);
            memcpy(slab, lap, sizeof lap);
        }
    }
}

void bwko_biharmonic_wk_scalar_cpu(const double *elem, double *qtens, const double *dvv, int nlev, int qsize, int nets, int nete)
{
    const size_t es = (size_t)16 * nlev * qsize;
    for (int ie = nets; ie <= nete; ++ie) element_cpu(elem + (size_t)ED * (ie - 1), qtens + es * (ie - 1), dvv, nlev, qsize);
}

void bwko_biharmonic_wk_scalar_cpu_omp(const double *elem, double *qtens, const double *dvv, int nlev, int qsize, int nets, int nete)
{
    const size_t es = (size_t)16 * nlev * qsize;
#pragma omp parallel for schedule(static)
    for (int ie = nets; ie <= nete; ++ie) element_cpu(elem + (size_t)ED * (ie - 1), qtens + es * (ie - 1), dvv, nlev, qsize);
}

int bwko_omp_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ================================================================= GPU1 order
 * F90:219-243 gradient, F90:247-278 divergence, F90:282-313 laplace, F90:317-360 driver.
 * One kchunk level at a time is enough: levels do not interact. */
static void laplace_gpu1_level(const double *s, const double *dvv, const double *e, double *lap)
{
    const double rrearth = bwko_rrearth();
    double g[32], og[32], vt[32];
    for (int j = 1; j <= NP; ++j)
        for (int i = 1; i <= NP; ++i) {
            double dsdx00 = 0.0, dsdy00 = 0.0;
            for (int l = 1; l <= NP; ++l) {
                dsdx00 = dsdx00 + DVV(dvv, l, i) * S2(s, l, j);
                dsdy00 = dsdy00 + DVV(dvv, l, j) * S2(s, i, l);
            }
            S3(g, i, j, 1) = (DINV(e, i, j, 1, 1) * dsdx00 + DINV(e, i, j, 2, 1) * dsdy00) * rrearth;
            S3(g, i, j, 2) = (DINV(e, i, j, 1, 2) * dsdx00 + DINV(e, i, j, 2, 2) * dsdy00) * rrearth;
        }
    memcpy(og, g, sizeof g);
    for (int j = 1; j <= NP; ++j)
        for (int i = 1; i <= NP; ++i) {
            S3(g, i, j, 1) = S3(og, i, j, 1) * TVISC(e, i, j, 1, 1) + S3(og, i, j, 2) * TVISC(e, i, j, 1, 2);
            S3(g, i, j, 2) = S3(og, i, j, 1) * TVISC(e, i, j, 2, 1) + S3(og, i, j, 2) * TVISC(e, i, j, 2, 2);
        }
    for (int j = 1; j <= NP; ++j)
        for (int i = 1; i <= NP; ++i) {
            S3(vt, i, j, 1) = (DINV(e, i, j, 1, 1) * S3(g, i, j, 1) + DINV(e, i, j, 1, 2) * S3(g, i, j, 2)) * SPMP(e, i, j) * rrearth;
            S3(vt, i, j, 2) = (DINV(e, i, j, 2, 1) * S3(g, i, j, 1) + DINV(e, i, j, 2, 2) * S3(g, i, j, 2)) * SPMP(e, i, j) * rrearth;
        }
    for (int j = 1; j <= NP; ++j)
        for (int i = 1; i <= NP; ++i) {
            double tmp = 0.0;
            for (int m = 1; m <= NP; ++m)
                tmp = tmp - S3(vt, m, j, 1) * DVV(dvv, i, m) - S3(vt, i, m, 2) * DVV(dvv, j, m);
            S2(lap, i, j) = tmp;
        }
}

void bwko_biharmonic_wk_scalar_gpu1(const double *elem, double *qtens, const double *dvv, int nlev, int qsize, int nets, int nete)
{
    const size_t es = (size_t)16 * nlev * qsize;
    const int nlev8 = (nlev / 8) * 8; /* F90:330: kc = 1 , nlev/kchunk */
    for (int ie = nets; ie <= nete; ++ie)
        for (int q = 0; q < qsize; ++q)
            for (int k = 0; k < nlev8; ++k) {
                double *slab = qtens + es * (ie - 1) + (size_t)16 * (k + (size_t)nlev * q);
                double s[16], lap[16];
                memcpy(s, slab, sizeof s);
                laplace_gpu1_level(s, dvv, elem + (size_t)ED * (ie - 1), lap);
                memcpy(slab, lap, sizeof lap);
            }
}

/* ================================================================= GPU2 order
 * F90:438-487 gradient kernel, F90:506-517 tensor loop, F90:381-434 divergence kernel,
 * F90:491-534 drivers.  grads is real global scratch as in the reference. */
void bwko_biharmonic_wk_scalar_gpu2(const double *elem, double *qtens, double *grads, const double *dvv,
                                    int nlev, int qsize, int nelemd, int nets, int nete)
{
    (void)nelemd;
    const double rrearth = bwko_rrearth();
    const int64_t len = (int64_t)nlev * qsize;
    /* kernel 1: gradient_sphere_openacc */
    for (int ie = nets; ie <= nete; ++ie) {
        const double *e = elem + (size_t)ED * (ie - 1);
        for (int64_t k = 0; k < len; ++k) {
            const double *s = qtens + (size_t)16 * (k + len * (ie - 1));
            double *ds = grads + (size_t)32 * (k + len * (ie - 1));
            for (int j = 1; j <= NP; ++j)
                for (int i = 1; i <= NP; ++i) {
                    double dsdx00 = 0.0, dsdy00 = 0.0;
                    for (int l = 1; l <= NP; ++l) {
                        dsdx00 = dsdx00 + DVV(dvv, l, i) * S2(s, l, j);
                        dsdy00 = dsdy00 + DVV(dvv, l, j) * S2(s, i, l);
                    }
                    S3(ds, i, j, 1) = (DINV(e, i, j, 1, 1) * dsdx00 + DINV(e, i, j, 2, 1) * dsdy00) * rrearth;
                    S3(ds, i, j, 2) = (DINV(e, i, j, 1, 2) * dsdx00 + DINV(e, i, j, 2, 2) * dsdy00) * rrearth;
                }
        }
    }
    /* kernel 2: tensor loop, sum(oldgrads(:)*tensorVisc(i,j,c,:)) */
    for (int ie = nets; ie <= nete; ++ie) {
        const double *e = elem + (size_t)ED * (ie - 1);
        for (int64_t k = 0; k < len; ++k) {
            double *g = grads + (size_t)32 * (k + len * (ie - 1));
            for (int j = 1; j <= NP; ++j)
                for (int i = 1; i <= NP; ++i) {
                    double o1 = S3(g, i, j, 1), o2 = S3(g, i, j, 2);
                    S3(g, i, j, 1) = o1 * TVISC(e, i, j, 1, 1) + o2 * TVISC(e, i, j, 1, 2);
                    S3(g, i, j, 2) = o1 * TVISC(e, i, j, 2, 1) + o2 * TVISC(e, i, j, 2, 2);
                }
        }
    }
    /* kernel 3: divergence_sphere_wk_openacc */
    for (int ie = nets; ie <= nete; ++ie) {
        const double *e = elem + (size_t)ED * (ie - 1);
        for (int64_t k = 0; k < len; ++k) {
            const double *v = grads + (size_t)32 * (k + len * (ie - 1));
            double *div = qtens + (size_t)16 * (k + len * (ie - 1));
            double vt[32];
            for (int j = 1; j <= NP; ++j)
                for (int i = 1; i <= NP; ++i) {
                    S3(vt, i, j, 1) = SPMP(e, i, j) * (DINV(e, i, j, 1, 1) * S3(v, i, j, 1) + DINV(e, i, j, 1, 2) * S3(v, i, j, 2));
                    S3(vt, i, j, 2) = SPMP(e, i, j) * (DINV(e, i, j, 2, 1) * S3(v, i, j, 1) + DINV(e, i, j, 2, 2) * S3(v, i, j, 2));
                }
            for (int j = 1; j <= NP; ++j)
                for (int i = 1; i <= NP; ++i) {
                    double tmp = 0.0;
                    for (int l = 1; l <= NP; ++l)
                        tmp = tmp - (S3(vt, l, j, 1) * DVV(dvv, i, l) + S3(vt, i, l, 2) * DVV(dvv, j, l));
                    S2(div, i, j) = tmp * rrearth;
                }
        }
    }
}

/* ------------------------------------------------------------------ F90:69-73 */
double bwko_compute_l2norm(const double *q, const double *qs, int64_t n)
{
    double num = 0.0, den = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        double d = q[i] - qs[i];
        num += d * d;
    }
    for (int64_t i = 0; i < n; ++i) den += qs[i] * qs[i];
    return sqrt(num / den);
}

double bwko_max_rel(const double *a, const double *b, int64_t n)
{
    double m = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        double d = fabs(a[i] - b[i]);
        double r = b[i] != 0.0 ? d / fabs(b[i]) : d;
        if (r > m || r != r) m = r != r ? INFINITY : r;
    }
    return m;
}

/* ===================== DSS + second sweep (upstream HOMME; PARITY UNPINNED) ===================== */

void bwko_cubed_sphere_gdof(int ne, int64_t *gdof)
{
    const int64_t L = 3 * (int64_t)ne, L1 = L + 1;
    for (int f = 0; f < 6; ++f)
        for (int je = 0; je < ne; ++je)
            for (int ie = 0; ie < ne; ++ie) {
                int64_t e = (int64_t)f * ne * ne + (int64_t)je * ne + ie;
                for (int j = 0; j < 4; ++j)
                    for (int i = 0; i < 4; ++i) {
                        int64_t A = 3 * ie + i, B = 3 * je + j, x, y, z;
                        switch (f) {
                        case 0: x = L; y = A; z = B; break;
                        case 1: x = L - A; y = L; z = B; break;
                        case 2: x = 0; y = L - A; z = B; break;
                        case 3: x = A; y = 0; z = B; break;
                        case 4: x = A; y = B; z = L; break;
                        default: x = A; y = B; z = 0; break;
                        }
                        gdof[16 * e + i + 4 * j] = x * L1 * L1 + y * L1 + z;
                    }
            }
}

struct bwko_dss {
    int64_t nelem, npts, ngroups;
    int64_t *start;  /* ngroups+1 */
    int64_t *member; /* npts, = 16*e + pt, ascending inside a group */
};

typedef struct { int64_t key, pt; } keypt_t;
static int cmp_keypt(const void *a, const void *b)
{
    const keypt_t *x = (const keypt_t *)a, *y = (const keypt_t *)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->pt < y->pt ? -1 : (x->pt > y->pt);
}

bwko_dss_t *bwko_dss_create(int64_t nelem, const int64_t *gdof)
{
    bwko_dss_t *d = (bwko_dss_t *)calloc(1, sizeof *d);
    d->nelem = nelem;
    d->npts = 16 * nelem;
    keypt_t *kp = (keypt_t *)malloc(sizeof(keypt_t) * d->npts);
    for (int64_t p = 0; p < d->npts; ++p) { kp[p].key = gdof[p]; kp[p].pt = p; }
    qsort(kp, d->npts, sizeof(keypt_t), cmp_keypt);
    d->member = (int64_t *)malloc(sizeof(int64_t) * d->npts);
    d->start = (int64_t *)malloc(sizeof(int64_t) * (d->npts + 1));
    int64_t g = 0;
    for (int64_t p = 0; p < d->npts; ++p) {
        if (p == 0 || kp[p].key != kp[p - 1].key) d->start[g++] = p;
        d->member[p] = kp[p].pt;
    }
    d->start[g] = d->npts;
    d->ngroups = g;
    free(kp);
    return d;
}

void bwko_dss_destroy(bwko_dss_t *d)
{
    if (!d) return;
    free(d->start);
    free(d->member);
    free(d);
}

int64_t bwko_dss_unique_nodes(const bwko_dss_t *d) { return d->ngroups; }

void bwko_dss_multiplicity_hist(const bwko_dss_t *d, int64_t *mult16)
{
    for (int m = 0; m < 16; ++m) mult16[m] = 0;
    for (int64_t g = 0; g < d->ngroups; ++g) {
        int64_t m = d->start[g + 1] - d->start[g];
        mult16[m < 15 ? m : 15]++;
    }
}

void bwko_dss_rspheremp(const bwko_dss_t *d, const double *elem, double *rsph)
{
    for (int64_t g = 0; g < d->ngroups; ++g) {
        double s = 0.0;
        for (int64_t t = d->start[g]; t < d->start[g + 1]; ++t) {
            int64_t p = d->member[t];
            s = (t == d->start[g]) ? elem[(size_t)ED * (p / 16) + 64 + (p % 16)] : s + elem[(size_t)ED * (p / 16) + 64 + (p % 16)];
        }
        for (int64_t t = d->start[g]; t < d->start[g + 1]; ++t) rsph[d->member[t]] = 1.0 / s;
    }
}

static void dss_apply_impl(const bwko_dss_t *d, double *qtens, int nlev, int qsize, int use_omp)
{
    const int64_t len = (int64_t)nlev * qsize;
    (void)use_omp;
#pragma omp parallel for schedule(static) if (use_omp)
    for (int64_t g = 0; g < d->ngroups; ++g) {
        const int64_t b = d->start[g], n = d->start[g + 1] - b;
        if (n < 2) continue;
        for (int64_t kq = 0; kq < len; ++kq) {
            double s = 0.0;
            for (int64_t t = 0; t < n; ++t) {
                int64_t p = d->member[b + t];
                double v = qtens[(size_t)16 * (kq + len * (p / 16)) + (p % 16)];
                s = t == 0 ? v : s + v;
            }
            for (int64_t t = 0; t < n; ++t) {
                int64_t p = d->member[b + t];
                qtens[(size_t)16 * (kq + len * (p / 16)) + (p % 16)] = s;
            }
        }
    }
}

void bwko_dss_apply(const bwko_dss_t *d, double *qtens, int nlev, int qsize) { dss_apply_impl(d, qtens, nlev, qsize, 0); }

static void element_sweep2(const double *e, const double *rs, double *qe, const double *dvv, int nlev, int qsize)
{
    for (int q = 0; q < qsize; ++q)
        for (int k = 0; k < nlev; ++k) {
            double *slab = qe + (size_t)16 * (k + (size_t)nlev * q);
            double s[16], lap[16];
            for (int t = 0; t < 16; ++t) s[t] = rs[t] * slab[t]; /* lap_p = rspheremp * qtens */
            laplace_sphere_wk_cpu(s, dvv, e, lap);
            memcpy(slab, lap, sizeof lap);
        }
}

void bwko_biharmonic_full(const bwko_dss_t *d, const double *elem, const double *rsph, double *qtens,
                          const double *dvv, int nlev, int qsize, int use_omp)
{
    const int nelem = (int)d->nelem;
    const size_t es = (size_t)16 * nlev * qsize;
    if (use_omp)
        bwko_biharmonic_wk_scalar_cpu_omp(elem, qtens, dvv, nlev, qsize, 1, nelem);
    else
        bwko_biharmonic_wk_scalar_cpu(elem, qtens, dvv, nlev, qsize, 1, nelem);
    dss_apply_impl(d, qtens, nlev, qsize, use_omp);
    if (use_omp) {
#pragma omp parallel for schedule(static)
        for (int ie = 0; ie < nelem; ++ie) element_sweep2(elem + (size_t)ED * ie, rsph + (size_t)16 * ie, qtens + es * ie, dvv, nlev, qsize);
    } else {
        for (int ie = 0; ie < nelem; ++ie) element_sweep2(elem + (size_t)ED * ie, rsph + (size_t)16 * ie, qtens + es * ie, dvv, nlev, qsize);
    }
}

/* ------------------------------------------------------------------------------------------
 * advance_hypervis_scalar -- NOT in /root/reference: the caller of
 * biharmonic_wk_scalar in upstream HOMME (prim_advection: advance_hypervis_scalar, limiter2d_zero).
 * Un-vendored, no pinned version, restated from its published structure; PARITY UNPINNED.
 * Per subcycle (dt = dt2 / subcycles):
 *     Qtens = dpdiss * Qdp / dp                           (hyperviscosity acts on the mixing ratio)
 *     Qtens = biharmonic_wk_scalar(Qtens)                 (Laplace -> DSS -> rspheremp -> Laplace, weak form)
 *     Qdp   = Qdp * spheremp - dt * nu_q * Qtens
 *     limiter2d_zero(Qdp(:,:,:,q))                        (optional: clip negatives level by level, keep the level's mass)
 *     Qdp   = DSS(Qdp) * rspheremp
 * Qdp(np,np,nlev,qsize,nelemd); dp, dpdiss (np,np,nlev,nelemd).
 */
static void limiter2d_zero_level(double *q16)
{
    double mass = 0.0;
    for (int t = 0; t < 16; ++t) mass = mass + q16[t];
    if (mass < 0.0)
        for (int t = 0; t < 16; ++t) q16[t] = -q16[t];
    double mass_added = 0.0;
    for (int t = 0; t < 16; ++t) {
        if (q16[t] < 0.0) q16[t] = 0.0;
        else mass_added = mass_added + q16[t];
    }
    if (mass_added > 0.0) {
        const double scale = fabs(mass) / mass_added;
        for (int t = 0; t < 16; ++t) q16[t] = q16[t] * scale;
    }
    if (mass < 0.0)
        for (int t = 0; t < 16; ++t) q16[t] = -q16[t];
}

void bwko_limiter2d_zero(double *q, int64_t nslabs)
{
    for (int64_t s = 0; s < nslabs; ++s) limiter2d_zero_level(q + 16 * s);
}

void bwko_advance_hypervis_scalar(const bwko_dss_t *d, const double *elem, const double *rsph, double *qdp, const double *dp,
                                  const double *dpdiss, const double *dvv, int nlev, int qsize, double dt2, double nu_q,
                                  int subcycles, int limiter)
{
    const int64_t nelem = d->nelem;
    const size_t es = (size_t)16 * nlev * qsize;
    double *qtens = (double *)malloc(sizeof(double) * es * (size_t)nelem);
    const double dt = dt2 / subcycles;
    for (int ic = 0; ic < subcycles; ++ic) {
        for (int64_t ie = 0; ie < nelem; ++ie)
            for (int q = 0; q < qsize; ++q)
                for (int k = 0; k < nlev; ++k)
                    for (int t = 0; t < 16; ++t) {
                        const size_t a = es * (size_t)ie + (size_t)16 * (k + (size_t)nlev * q) + t;
                        const size_t b = (size_t)16 * nlev * (size_t)ie + (size_t)16 * k + t;
                        qtens[a] = dpdiss[b] * qdp[a] / dp[b];
                    }
        bwko_biharmonic_full(d, elem, rsph, qtens, dvv, nlev, qsize, 0);
        for (int64_t ie = 0; ie < nelem; ++ie) {
            const double *e = elem + (size_t)ED * ie;
            for (int q = 0; q < qsize; ++q)
                for (int k = 0; k < nlev; ++k) {
                    double *slab = qdp + es * (size_t)ie + (size_t)16 * (k + (size_t)nlev * q);
                    const double *ts = qtens + es * (size_t)ie + (size_t)16 * (k + (size_t)nlev * q);
                    for (int t = 0; t < 16; ++t) slab[t] = slab[t] * e[64 + t] - dt * nu_q * ts[t];
                    if (limiter) limiter2d_zero_level(slab);
                }
        }
        dss_apply_impl(d, qdp, nlev, qsize, 0);
        for (int64_t ie = 0; ie < nelem; ++ie)
            for (size_t s = 0; s < (size_t)nlev * qsize; ++s)
                for (int t = 0; t < 16; ++t) qdp[es * (size_t)ie + 16 * s + t] = rsph[(size_t)16 * ie + t] * qdp[es * (size_t)ie + 16 * s + t];
    }
    free(qtens);
}

/* ------------------------------------------------------------------------------------------
 * biharmonic_wk_dp3d / vlaplace_sphere_wk -- NOT in /root/reference: the momentum /
 * temperature / layer-thickness counterpart of biharmonic_wk_scalar in upstream HOMME (viscosity_mod:
 * biharmonic_wk_dp3d; derivative_mod: vlaplace_sphere_wk, vlaplace_sphere_wk_cartesian).  Un-vendored, no pinned
 * version, restated from its published structure; PARITY UNPINNED (pinned to analytic eigenvalues by tests/test_dp3d.py).
 *
 *   ttens  = laplace_sphere_wk(T),  dptens = laplace_sphere_wk(dp3d),  vtens = vlaplace_sphere_wk(v)      [var_coef = var_coef1]
 *   DSS of all four fields (one edge buffer: T at 0, v at nlev, dp3d at 3 nlev), * rspheremp
 *   the same three operators again                                                                          [var_coef = .true.]
 *
 * vlaplace_sphere_wk in its tensor-hyperviscosity form (the branch the reference's laplace_sphere_wk, F90:164-182,
 * belongs to) is the Cartesian one: the lat-lon vector goes to three Cartesian components with vec_sphere2cart(np,np,3,2),
 * each component gets laplace_sphere_wk, the result comes back with the transpose, and 2*spheremp*v*rrearth^2 is added so
 * that rigid rotation is not damped ("undamprrcart").  var_coef = 0 applies the operator without tensorVisc (upstream
 * passes var_coef1 = .false. on the first application when hypervis_scaling > 0); upstream's other branch (the
 * contravariant grad-div / curl-vorticity form) is not restated here.
 * Layouts: ttens, dptens (np,np,nlev,nelemd); vtens (np,np,2,nlev,nelemd); vec_sphere2cart (np,np,3,2,nelemd).
 */
static void laplace_sphere_wk_vc(const double *s, const double *dvv, const double *e, int var_coef, double *laplace)
{
    if (var_coef) {
        laplace_sphere_wk_cpu(s, dvv, e, laplace);
        return;
    }
    double grads[32];
    gradient_sphere_cpu(s, dvv, e, grads);
    divergence_sphere_wk_cpu(grads, dvv, e, laplace);
}

static void vlaplace_cartesian_cpu(const double *v, const double *dvv, const double *e, const double *v2c, int var_coef, int undamp, double *lap)
{
    const double rrearth = bwko_rrearth();
    double cart[3][16], lc[3][16];
    for (int c = 0; c < 3; ++c)
        for (int t = 0; t < 16; ++t) cart[c][t] = v2c[t + 16 * c] * v[t] + v2c[t + 16 * c + 48] * v[16 + t];
    for (int c = 0; c < 3; ++c) laplace_sphere_wk_vc(cart[c], dvv, e, var_coef, lc[c]);
    for (int m = 0; m < 2; ++m)
        for (int t = 0; t < 16; ++t)
            lap[16 * m + t] = lc[0][t] * v2c[t + 48 * m] + lc[1][t] * v2c[t + 16 + 48 * m] + lc[2][t] * v2c[t + 32 + 48 * m];
    if (undamp)
        for (int m = 0; m < 2; ++m)
            for (int t = 0; t < 16; ++t) lap[16 * m + t] = lap[16 * m + t] + 2 * e[64 + t] * v[16 * m + t] * (rrearth * rrearth);
}

void bwko_vlaplace_sphere_wk(const double *elem, const double *v2c, double *v, const double *dvv, int nlev, int64_t nelem, int var_coef,
                             int undamp)
{
    for (int64_t ie = 0; ie < nelem; ++ie)
        for (int k = 0; k < nlev; ++k) {
            double *slab = v + ((size_t)nlev * ie + k) * 32;
            double in[32], out[32];
            memcpy(in, slab, sizeof in);
            vlaplace_cartesian_cpu(in, dvv, elem + (size_t)ED * ie, v2c + (size_t)96 * ie, var_coef, undamp, out);
            memcpy(slab, out, sizeof out);
        }
}

static void dp3d_sweep(const double *elem, const double *v2c, const double *rsph, double *dptens, double *ttens, double *vtens, const double *dvv,
                       int nlev, int64_t nelem, int var_coef, int undamp)
{
    for (int64_t ie = 0; ie < nelem; ++ie) {
        const double *e = elem + (size_t)ED * ie;
        for (int k = 0; k < nlev; ++k) {
            double *sl[2] = {ttens + ((size_t)nlev * ie + k) * 16, dptens + ((size_t)nlev * ie + k) * 16};
            for (int f = 0; f < 2; ++f) {
                double s[16], lap[16];
                for (int t = 0; t < 16; ++t) s[t] = rsph ? rsph[(size_t)16 * ie + t] * sl[f][t] : sl[f][t];
                laplace_sphere_wk_vc(s, dvv, e, var_coef, lap);
                memcpy(sl[f], lap, sizeof lap);
            }
            double *vs = vtens + ((size_t)nlev * ie + k) * 32;
            double in[32], out[32];
            for (int m = 0; m < 2; ++m)
                for (int t = 0; t < 16; ++t) in[16 * m + t] = rsph ? rsph[(size_t)16 * ie + t] * vs[16 * m + t] : vs[16 * m + t];
            vlaplace_cartesian_cpu(in, dvv, e, v2c + (size_t)96 * ie, var_coef, undamp, out);
            memcpy(vs, out, sizeof out);
        }
    }
}

void bwko_biharmonic_wk_dp3d(const bwko_dss_t *d, const double *elem, const double *v2c, const double *rsph, double *dptens, double *ttens,
                             double *vtens, const double *dvv, int nlev, int var_coef1, int undamp)
{
    const int64_t nelem = d->nelem;
    dp3d_sweep(elem, v2c, NULL, dptens, ttens, vtens, dvv, nlev, nelem, var_coef1, undamp);
    dss_apply_impl(d, ttens, nlev, 1, 0);
    dss_apply_impl(d, vtens, 2 * nlev, 1, 0);
    dss_apply_impl(d, dptens, nlev, 1, 0);
    dp3d_sweep(elem, v2c, rsph, dptens, ttens, vtens, dvv, nlev, nelem, 1, undamp);
}
