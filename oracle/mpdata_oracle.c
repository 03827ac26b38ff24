/* TEST INFRASTRUCTURE ONLY -- CPU oracle for the mmf-mpdata-tracer mini-app.
 *
 * Plain-C restatement of subroutine advect_scalar2D_cpu
 * (mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90:477-640, cited below as A:<line>):
 * positive-definite monotonic (MPDATA) advection of one scalar on nslices independent x-z planes,
 * slice index fastest ("pushncols").  Loop order, operation order and the in-place updates follow
 * the Fortran; build with -ffp-contract=off.  Only tests/, __graft_entry__ and bench/profile CPU
 * legs may load this; the product (libmpdata_b200.so) never does.
 *
 * PARITY: pinned on the reference's own source, mechanically translated.  The reference stores no outputs
 * for this kernel and draws its inputs from the compiler's random_number (A:645-665); its own check is
 * "relative L1 error of f and flux against the CPU subroutine" (A:679-683): exactly 0 for gnu.7.1.0, up to
 * 1e-4 for other builds under results/ (tests/golden/mpdata_results.json).  There is no Fortran compiler in
 * this image, so `make -C oracle ref` translates A:477-642 statement by statement to C (oracle/f90_to_c.py,
 * a Fortran-subset translator that refuses what it does not know) and compiles it into oracle/_ref;
 * tests/test_mpdata_ref.py requires THIS file to equal that build bit for bit (7 shapes, both velocity
 * signs) and reproduces the reference's acceptance test on the translated OpenACC forms (exactly 0).
 * Not the output of a Fortran compiler -- see oracle/ref_mpdata_py.py for what the claim is.
 *
 * Array shapes (Fortran, column-major; A:14-36), nzm = nz-1:
 *   f   (nslices, -2:nx+3, 1, nzm)    u (nslices, -1:nx+3, 1, nzm)    w (nslices, -1:nx+2, 1, nz)
 *   rho (nslices, nzm)  rhow (nslices, nz)  adz (nslices, nzm)  flux (nslices, nz)
 * `eps = 1.e-10` is a default-real literal assigned to a double (A:509): the value is float32(1e-10).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define F(sl, i, k) f[(sl) + ns * (size_t)(((i) + 2) + FX * ((k)-1))]
#define U(sl, i, k) u[(sl) + ns * (size_t)(((i) + 1) + UX * ((k)-1))]
#define W(sl, i, k) w[(sl) + ns * (size_t)(((i) + 1) + WX * ((k)-1))]
#define UUU(sl, i, k) uuu[(sl) + ns * (size_t)(((i) + 1) + UX * ((k)-1))]
#define WWW(sl, i, k) www[(sl) + ns * (size_t)(((i) + 1) + WX * ((k)-1))]
#define MX(sl, i, k) mx[(sl) + ns * (size_t)((i) + MXW * ((k)-1))]
#define MN(sl, i, k) mn[(sl) + ns * (size_t)((i) + MXW * ((k)-1))]
#define S2(a, sl, k) a[(sl) + ns * (size_t)((k)-1)]

/* Timing build only (-fopenmp): every level loop of the step is independent across k (the reference's OpenACC forms
 * collapse the same loops, A:112-227), so threads take levels; no reduction crosses levels, the bits do not change. */
#define OMP_K _Pragma("omp parallel for schedule(static) private(i, sl, kb, kc, ib, ic)")

/* Fortran MAX/MIN as compare-select (what gfortran -O3 emits; libm fmax/fmin are calls and 2x slower) */
static inline double dmax(double a, double b) { return a > b ? a : b; }
static inline double dmin(double a, double b) { return a < b ? a : b; }
/* statement functions, A:500-503 */
static inline double andiff(double x1, double x2, double a, double b) { return (fabs(a) - a * a * b) * 0.5 * (x2 - x1); }
static inline double across(double x1, double a1, double a2) { return 0.03125 * a1 * a2 * x1; }
static inline double pp(double y) { return dmax(0.0, y); }
static inline double pn(double y) { return -dmin(0.0, y); }
static inline double max5(double a, double b, double c, double d, double e) { return dmax(dmax(dmax(dmax(a, b), c), d), e); }
static inline double min5(double a, double b, double c, double d, double e) { return dmin(dmin(dmin(dmin(a, b), c), d), e); }
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

/* Returns 0, or 1 on bad sizes / allocation failure.  f is updated in place, flux(:,1:nzm) is
 * written, flux(:,nz) is left untouched exactly as the reference leaves it. */
int mpdo_advect_scalar2D_cpu(int ns, int nx, int nz, double *f, const double *u, const double *w, const double *rho,
                             const double *rhow, const double *adz, double *flux) {
    if (ns < 1 || nx < 1 || nz < 3) return 1;
    const int nzm = nz - 1, nxp1 = nx + 1, nxp2 = nx + 2, nxp3 = nx + 3;
    const int FX = nx + 6, UX = nx + 5, WX = nx + 4, MXW = nx + 2;
    const double eps = (double)1.e-10f; /* A:509 */
    double *mx = malloc(sizeof(double) * ns * MXW * nzm), *mn = malloc(sizeof(double) * ns * MXW * nzm);
    double *uuu = malloc(sizeof(double) * ns * UX * nzm), *www = malloc(sizeof(double) * ns * WX * nz);
    double *iadz = malloc(sizeof(double) * ns * nzm), *irho = malloc(sizeof(double) * ns * nzm);
    double *irhow = malloc(sizeof(double) * ns * nzm);
    if (!mx || !mn || !uuu || !www || !iadz || !irho || !irhow) return 1;
    int i, k, sl, kc, kb, ib, ic;

    for (i = -1; i <= nxp2; ++i) /* A:511 */
        for (sl = 0; sl < ns; ++sl) WWW(sl, i, nz) = 0.0;

    OMP_K
    for (k = 1; k <= nzm; ++k) { /* A:513-526 */
        kc = imin(nzm, k + 1);
        kb = imax(1, k - 1);
        for (i = 0; i <= nxp1; ++i)
            for (sl = 0; sl < ns; ++sl) {
                ib = i - 1;
                ic = i + 1;
                MX(sl, i, k) = max5(F(sl, ib, k), F(sl, ic, k), F(sl, i, kb), F(sl, i, kc), F(sl, i, k));
                MN(sl, i, k) = min5(F(sl, ib, k), F(sl, ic, k), F(sl, i, kb), F(sl, i, kc), F(sl, i, k));
            }
    }

    OMP_K
    for (k = 1; k <= nzm; ++k) { /* A:528-548 */
        kb = imax(1, k - 1);
        for (i = -1; i <= nxp3; ++i)
            for (sl = 0; sl < ns; ++sl)
                UUU(sl, i, k) = dmax(0.0, U(sl, i, k)) * F(sl, i - 1, k) + dmin(0.0, U(sl, i, k)) * F(sl, i, k);
        for (i = -1; i <= nxp2; ++i)
            for (sl = 0; sl < ns; ++sl)
                WWW(sl, i, k) = dmax(0.0, W(sl, i, k)) * F(sl, i, kb) + dmin(0.0, W(sl, i, k)) * F(sl, i, k);
        for (sl = 0; sl < ns; ++sl) S2(flux, sl, k) = 0.0;
        for (i = 1; i <= nx; ++i)
            for (sl = 0; sl < ns; ++sl) S2(flux, sl, k) = S2(flux, sl, k) + WWW(sl, i, k);
    }

    OMP_K
    for (k = 1; k <= nzm; ++k) { /* A:550-560 */
        for (sl = 0; sl < ns; ++sl) {
            S2(irho, sl, k) = 1.0 / S2(rho, sl, k);
            S2(iadz, sl, k) = 1.0 / S2(adz, sl, k);
        }
        for (i = -1; i <= nxp2; ++i)
            for (sl = 0; sl < ns; ++sl)
                F(sl, i, k) = F(sl, i, k) - (UUU(sl, i + 1, k) - UUU(sl, i, k) + (WWW(sl, i, k + 1) - WWW(sl, i, k)) * S2(iadz, sl, k)) * S2(irho, sl, k);
    }

    OMP_K
    for (k = 1; k <= nzm; ++k) { /* A:561-585 */
        kc = imin(nzm, k + 1);
        kb = imax(1, k - 1);
        for (sl = 0; sl < ns; ++sl) S2(irhow, sl, k) = 1.0 / (S2(rhow, sl, k) * S2(adz, sl, k));
        for (i = 0; i <= nxp2; ++i)
            for (sl = 0; sl < ns; ++sl) {
                double dd = 2.0 / (double)(kc - kb) / S2(adz, sl, k);
                ib = i - 1;
                UUU(sl, i, k) = andiff(F(sl, ib, k), F(sl, i, k), U(sl, i, k), S2(irho, sl, k)) -
                                across(dd * (F(sl, ib, kc) + F(sl, i, kc) - F(sl, ib, kb) - F(sl, i, kb)), U(sl, i, k),
                                       W(sl, ib, k) + W(sl, ib, kc) + W(sl, i, k) + W(sl, i, kc)) * S2(irho, sl, k);
            }
        for (i = 0; i <= nxp1; ++i)
            for (sl = 0; sl < ns; ++sl) {
                ib = i - 1;
                ic = i + 1;
                WWW(sl, i, k) = andiff(F(sl, i, kb), F(sl, i, k), W(sl, i, k), S2(irhow, sl, k)) -
                                across(F(sl, ic, kb) + F(sl, ic, k) - F(sl, ib, kb) - F(sl, ib, k), W(sl, i, k),
                                       U(sl, i, kb) + U(sl, i, k) + U(sl, ic, k) + U(sl, ic, kb)) * S2(irho, sl, k);
            }
    }
    for (i = -1; i <= nxp2; ++i) /* A:586 */
        for (sl = 0; sl < ns; ++sl) WWW(sl, i, 1) = 0.0;

    OMP_K
    for (k = 1; k <= nzm; ++k) { /* A:589-600 */
        kc = imin(nzm, k + 1);
        kb = imax(1, k - 1);
        for (i = 0; i <= nxp1; ++i)
            for (sl = 0; sl < ns; ++sl) {
                ib = i - 1;
                ic = i + 1;
                MX(sl, i, k) = dmax(max5(F(sl, ib, k), F(sl, ic, k), F(sl, i, kb), F(sl, i, kc), F(sl, i, k)), MX(sl, i, k));
                MN(sl, i, k) = dmin(min5(F(sl, ib, k), F(sl, ic, k), F(sl, i, kb), F(sl, i, kc), F(sl, i, k)), MN(sl, i, k));
            }
    }
    OMP_K
    for (k = 1; k <= nzm; ++k) { /* A:601-612 */
        kc = imin(nzm, k + 1);
        for (i = 0; i <= nxp1; ++i)
            for (sl = 0; sl < ns; ++sl) {
                ic = i + 1;
                MX(sl, i, k) = S2(rho, sl, k) * (MX(sl, i, k) - F(sl, i, k)) /
                               (pn(UUU(sl, ic, k)) + pp(UUU(sl, i, k)) + S2(iadz, sl, k) * (pn(WWW(sl, i, kc)) + pp(WWW(sl, i, k))) + eps);
                MN(sl, i, k) = S2(rho, sl, k) * (F(sl, i, k) - MN(sl, i, k)) /
                               (pp(UUU(sl, ic, k)) + pn(UUU(sl, i, k)) + S2(iadz, sl, k) * (pp(WWW(sl, i, kc)) + pn(WWW(sl, i, k))) + eps);
            }
    }
    OMP_K
    for (k = 1; k <= nzm; ++k) { /* A:613-627 */
        kb = imax(1, k - 1);
        for (i = 1; i <= nxp1; ++i)
            for (sl = 0; sl < ns; ++sl) {
                ib = i - 1;
                UUU(sl, i, k) = pp(UUU(sl, i, k)) * dmin(dmin(1.0, MX(sl, i, k)), MN(sl, ib, k)) -
                                pn(UUU(sl, i, k)) * dmin(dmin(1.0, MX(sl, ib, k)), MN(sl, i, k));
            }
        for (i = 1; i <= nx; ++i)
            for (sl = 0; sl < ns; ++sl) {
                WWW(sl, i, k) = pp(WWW(sl, i, k)) * dmin(dmin(1.0, MX(sl, i, k)), MN(sl, i, kb)) -
                                pn(WWW(sl, i, k)) * dmin(dmin(1.0, MX(sl, i, kb)), MN(sl, i, k));
                S2(flux, sl, k) = S2(flux, sl, k) + WWW(sl, i, k);
            }
    }

    OMP_K
    for (k = 1; k <= nzm; ++k) /* A:630-637 */
        for (i = 1; i <= nx; ++i)
            for (sl = 0; sl < ns; ++sl)
                F(sl, i, k) = dmax(0.0, F(sl, i, k) - (UUU(sl, i + 1, k) - UUU(sl, i, k) + (WWW(sl, i, k + 1) - WWW(sl, i, k)) * S2(iadz, sl, k)) * S2(irho, sl, k));

    free(mx); free(mn); free(uuu); free(www); free(iadz); free(irho); free(irhow);
    return 0;
}

/* The reference's acceptance quantity (A:679-683): sum|a-b| / sum|b|. */
double mpdo_relative_l1(const double *a, const double *b, size_t n) {
    double num = 0.0, den = 0.0;
    for (size_t j = 0; j < n; ++j) {
        num += fabs(a[j] - b[j]);
        den += fabs(b[j]);
    }
    return num / den;
}

/* Timing leg (builder extension): the slices are independent, so blocks of slices can be run on
 * separate threads by calling the restatement on sub-ranges -- but the slice index is the fastest
 * one in memory, so each thread gathers its block into a private contiguous copy first. */
int mpdo_advect_scalar2D_cpu_blocks(int ns, int nx, int nz, double *f, const double *u, const double *w, const double *rho,
                                    const double *rhow, const double *adz, double *flux, int block) {
    if (block < 1) return 1;
    const int nzm = nz - 1;
    const size_t pf = (size_t)(nx + 6) * nzm, pu = (size_t)(nx + 5) * nzm, pw = (size_t)(nx + 4) * nz;
    int nb = (ns + block - 1) / block, rc = 0;
#pragma omp parallel for schedule(dynamic) reduction(| : rc)
    for (int b = 0; b < nb; ++b) {
        int s0 = b * block, n = imin(block, ns - s0);
        double *lf = malloc(8 * n * pf), *lu = malloc(8 * n * pu), *lw = malloc(8 * n * pw);
        double *lr = malloc(8 * n * nzm), *lrw = malloc(8 * n * nz), *la = malloc(8 * n * nzm), *lfl = malloc(8 * n * nz);
#define GATHER(dst, src, planes) for (size_t p = 0; p < (planes); ++p) memcpy(dst + p * n, src + p * ns + s0, 8 * n)
#define SCATTER(dst, src, planes) for (size_t p = 0; p < (planes); ++p) memcpy(dst + p * ns + s0, src + p * n, 8 * n)
        GATHER(lf, f, pf); GATHER(lu, u, pu); GATHER(lw, w, pw);
        GATHER(lr, rho, (size_t)nzm); GATHER(lrw, rhow, (size_t)nz); GATHER(la, adz, (size_t)nzm); GATHER(lfl, flux, (size_t)nz);
        rc |= mpdo_advect_scalar2D_cpu(n, nx, nz, lf, lu, lw, lr, lrw, la, lfl);
        SCATTER(f, lf, pf); SCATTER(flux, lfl, (size_t)nz);
        free(lf); free(lu); free(lw); free(lr); free(lrw); free(la); free(lfl);
    }
    return rc;
}
