/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of the MPAS-Ocean high-order advection flux
 * loop of the reference mini-app nested_loops/ (cited as nested.F90:<line>, cke_impl1.cpp:<line>).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may link or call this.
 *
 * PARITY: pinned on the reference compiled here.  The reference stores no outputs for this kernel (no
 * results/ directory, its inputs come from the compiler's random_number), but its plain-C++ form
 * nested_loops/nested_yakl_c.cpp:142-306 compiles in this image: `make -C oracle ref` builds it where it
 * lies into oracle/_ref/libnested_yakl_ref.so (oracle/ref_yakl_py.py) and tests/test_zz_nested.py requires
 * the forms below to equal it bit for bit.  The reference's own acceptance test -- every loop form must
 * agree with the "original CPU" form to errTol = 1e-10 relative (nested_vars.F90:33, nested.F90:267-279) --
 * is reproduced there too, plus an independent numpy restatement.
 *
 * Arrays keep the Fortran layouts: x(nvldim, n) column-major == C x[n][nvldim] (level fastest),
 * advCellsForEdge(nAdv, nEdges) == C [nEdges][nAdv]; cell and level indices are 1-based as in
 * the Fortran caller.  Compile with -ffp-contract=off: a*b*c is evaluated left to right as
 * Fortran does, no FMA contraction.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

typedef struct {
    int nEdges, nCells, nVertLevels, nvldim, nAdv;
    const int *nAdvCellsForEdge, *minLevelCell, *maxLevelCell, *advCellsForEdge;
    const double *tracerCur, *cellMask, *normalThicknessFlux, *advMaskHighOrder;
    const double *advCoefs, *advCoefs3rd;
    double coef3rdOrder;
} ckeo_problem_t;

/* Fortran sign(1.0, x): magnitude 1 with the sign bit of x (gfortran: -0.0 -> -1). */
static double f_sign1(double x) { return copysign(1.0, x); }

/* The reference answer refFlx: "original CPU" loop form, nested.F90:119-157.
 * edge -> (per-level factors) -> contributing cell -> levels kmin..kmax of that cell. */
void ckeo_original_cpu(const ckeo_problem_t *p, double *flx) {
    const int nv = p->nvldim;
    double *wgt = (double *)malloc(sizeof(double) * (size_t)nv);
    double *sgn = (double *)malloc(sizeof(double) * (size_t)nv);
    for (int e = 0; e < p->nEdges; ++e) {
        const double *ntf = p->normalThicknessFlux + (size_t)e * nv;
        const double *msk = p->advMaskHighOrder + (size_t)e * nv;
        double *out = flx + (size_t)e * nv;
        for (int k = 0; k < p->nVertLevels; ++k) { /* nested.F90:130-136 */
            wgt[k] = ntf[k] * msk[k];
            sgn[k] = f_sign1(ntf[k]);
            out[k] = 0.0;
        }
        for (int i = 0; i < p->nAdvCellsForEdge[e]; ++i) { /* nested.F90:140-155 */
            const int c = p->advCellsForEdge[(size_t)e * p->nAdv + i] - 1;
            const int kmin = p->minLevelCell[c], kmax = p->maxLevelCell[c];
            const double coef1 = p->advCoefs[(size_t)e * p->nAdv + i];
            const double coef3 = p->advCoefs3rd[(size_t)e * p->nAdv + i] * p->coef3rdOrder;
            const double *tr = p->tracerCur + (size_t)c * nv;
            for (int k = kmin - 1; k <= kmax - 1; ++k) {
                const double t = tr[k] * wgt[k];
                out[k] = out[k] + t * (coef1 + coef3 * sgn[k]);
            }
        }
    }
    free(wgt);
    free(sgn);
}

/* "GPU optimized" loop form, nested.F90:606-630: collapse(edge, level), the level range of the
 * cell replaced by a multiplication with cellMask. */
void ckeo_gpu_optimized(const ckeo_problem_t *p, double *flx) {
    const int nv = p->nvldim;
    for (int e = 0; e < p->nEdges; ++e)
        for (int k = 0; k < p->nVertLevels; ++k) {
            const double ntf = p->normalThicknessFlux[(size_t)e * nv + k];
            const double coef2 = ntf * p->advMaskHighOrder[(size_t)e * nv + k];
            const double csgn = f_sign1(ntf);
            double edgeFlx = 0.0;
            for (int i = 0; i < p->nAdvCellsForEdge[e]; ++i) {
                const int c = p->advCellsForEdge[(size_t)e * p->nAdv + i] - 1;
                const double coef1 = p->advCoefs[(size_t)e * p->nAdv + i];
                const double coef3 = p->advCoefs3rd[(size_t)e * p->nAdv + i] * p->coef3rdOrder;
                const double t = p->tracerCur[(size_t)c * nv + k] * p->cellMask[(size_t)c * nv + k];
                edgeFlx = edgeFlx + (t * coef2) * (coef1 + coef3 * csgn);
            }
            flx[(size_t)e * nv + k] = edgeFlx;
        }
}

/* One iteration of C++/Kokkos implementation 1, cke_impl1.cpp:11-37 (and the k-tiled Fortran form
 * nested.F90:307-349): pass 1 masks the tracer IN PLACE, pass 2 is the collapsed loop with
 * csgn = ntf < 0 ? -1 : 1 over all nvldim levels (whole packs, cke_impl.hpp:96-98). */
void ckeo_impl1(const ckeo_problem_t *p, double *tracerCur, double *flx) {
    const int nv = p->nvldim;
    for (size_t n = 0; n < (size_t)p->nCells * nv; ++n) tracerCur[n] = tracerCur[n] * p->cellMask[n];
    for (int e = 0; e < p->nEdges; ++e)
        for (int k = 0; k < nv; ++k) {
            const double ntf = p->normalThicknessFlux[(size_t)e * nv + k];
            const double coef2 = ntf * p->advMaskHighOrder[(size_t)e * nv + k];
            const double csgn = ntf < 0 ? -1.0 : 1.0;
            double edgeFlx = 0.0;
            for (int i = 0; i < p->nAdvCellsForEdge[e]; ++i) {
                const int c = p->advCellsForEdge[(size_t)e * p->nAdv + i] - 1;
                const double coef1 = p->advCoefs[(size_t)e * p->nAdv + i];
                const double coef3 = p->advCoefs3rd[(size_t)e * p->nAdv + i] * p->coef3rdOrder;
                edgeFlx = edgeFlx + (tracerCur[(size_t)c * nv + k] * coef2) * (coef1 + coef3 * csgn);
            }
            flx[(size_t)e * nv + k] = edgeFlx;
        }
}

/* One iteration of C++/Kokkos implementation 2 at pack size 1, cke_impl2.cpp:27-68: per edge,
 * level factors first, then each contributing cell adds tracer*mask*wgt*(coef1+coef3*sgn) on
 * its own level range only; levels no cell reaches stay 0. */
void ckeo_impl2(const ckeo_problem_t *p, const double *tracerCur, double *flx) {
    const int nv = p->nvldim;
    double *wgt = (double *)malloc(sizeof(double) * (size_t)nv);
    double *sgn = (double *)malloc(sizeof(double) * (size_t)nv);
    for (int e = 0; e < p->nEdges; ++e) {
        double *out = flx + (size_t)e * nv;
        for (int k = 0; k < nv; ++k) {
            const double ntf = p->normalThicknessFlux[(size_t)e * nv + k];
            wgt[k] = ntf * p->advMaskHighOrder[(size_t)e * nv + k];
            sgn[k] = ntf < 0 ? -1.0 : 1.0;
            out[k] = 0.0;
        }
        for (int i = 0; i < p->nAdvCellsForEdge[e]; ++i) {
            const int c = p->advCellsForEdge[(size_t)e * p->nAdv + i] - 1;
            const double coef1 = p->advCoefs[(size_t)e * p->nAdv + i];
            const double coef3 = p->advCoefs3rd[(size_t)e * p->nAdv + i] * p->coef3rdOrder;
            for (int k = p->minLevelCell[c] - 1; k <= p->maxLevelCell[c] - 1; ++k) {
                const double t = tracerCur[(size_t)c * nv + k] * p->cellMask[(size_t)c * nv + k];
                out[k] = out[k] + (t * wgt[k]) * (coef1 + coef3 * sgn[k]);
            }
        }
    }
    free(wgt);
    free(sgn);
}

/* The reference's acceptance test, nested.F90:267-279: number of points whose relative error
 * against refFlx exceeds errTol (absolute error where refFlx == 0); NaNs count as failures
 * (nested.F90:425).  *max_rel receives the largest error seen. */
int64_t ckeo_check(const double *flx, const double *refFlx, int nEdges, int nVertLevels, int nvldim, double errTol,
                   double *max_rel) {
    int64_t bad = 0;
    double worst = 0.0;
    for (int e = 0; e < nEdges; ++e)
        for (int k = 0; k < nVertLevels; ++k) {
            const double refVal = refFlx[(size_t)e * nvldim + k];
            double relErr = fabs(flx[(size_t)e * nvldim + k] - refVal);
            if (refVal != 0.0) relErr = relErr / fabs(refVal);
            if (isnan(relErr) || relErr > errTol) ++bad;
            if (relErr > worst || isnan(relErr)) worst = relErr;
        }
    if (max_rel) *max_rel = worst;
    return bad;
}

/* Timing leg for bench/profiles (cpu_baseline kind "port"): iters passes of the original CPU form,
 * OpenMP over edges when built with -fopenmp (builder extension; the reference loop is serial). */
void ckeo_original_cpu_omp(const ckeo_problem_t *p, double *flx, int iters) {
    const int nv = p->nvldim;
    for (int it = 0; it < iters; ++it) {
#pragma omp parallel
        {
            double *wgt = (double *)malloc(sizeof(double) * (size_t)nv * 2);
            double *sgn = wgt + nv;
#pragma omp for schedule(static)
            for (int e = 0; e < p->nEdges; ++e) {
                const double *ntf = p->normalThicknessFlux + (size_t)e * nv;
                const double *msk = p->advMaskHighOrder + (size_t)e * nv;
                double *out = flx + (size_t)e * nv;
                for (int k = 0; k < p->nVertLevels; ++k) {
                    wgt[k] = ntf[k] * msk[k];
                    sgn[k] = f_sign1(ntf[k]);
                    out[k] = 0.0;
                }
                for (int i = 0; i < p->nAdvCellsForEdge[e]; ++i) {
                    const int c = p->advCellsForEdge[(size_t)e * p->nAdv + i] - 1;
                    const double coef1 = p->advCoefs[(size_t)e * p->nAdv + i];
                    const double coef3 = p->advCoefs3rd[(size_t)e * p->nAdv + i] * p->coef3rdOrder;
                    const double *tr = p->tracerCur + (size_t)c * nv;
                    for (int k = p->minLevelCell[c] - 1; k <= p->maxLevelCell[c] - 1; ++k)
                        out[k] = out[k] + (tr[k] * wgt[k]) * (coef1 + coef3 * sgn[k]);
                }
            }
            free(wgt);
        }
    }
}
