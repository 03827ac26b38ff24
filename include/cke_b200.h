/*
 * cke_b200.h -- C ABI of libcke_b200.so: the B200 (sm_100a) implementation of the MPAS-Ocean
 * high-order advection flux loop of the reference mini-app nested_loops/ (SURVEY 8(f)4).
 *
 * The reference already has a C boundary for this kernel: nested_loops/cke.hpp:7-23, bound from
 * Fortran by nested_loops/cke_mod.F90:4-50 and called by nested_loops/nested.F90:391-433 when
 * built with -DUSE_CKE.  The first block below repeats that interface name for name and argument
 * for argument, so the reference's own cke_mod.F90 binds to this library unchanged: link
 * libcke_b200.so in place of cke.o + cke_impl1.o + cke_impl2.o + Kokkos/EKAT.
 *
 * Array layouts are the Fortran caller's: x(nvldim, n) column-major == C x[n][nvldim] (level
 * fastest), advCellsForEdge(nAdv, nEdges) == C [nEdges][nAdv]; cell and level indices 1-based
 * (cke.cpp:104-107 subtracts 1; so does this library, on the device).
 *
 * Error behaviour: the reference's entry points return void and assert (cke.cpp:142, cke_impl1.cpp:44).
 * Here a failure (no sm_100 device, allocation, launch) prints the message to stderr and
 * abort()s, unless cke_b200_set_abort_on_error(0) was called, in which case the call returns and
 * cke_b200_last_error() is non-NULL until the next cke_init.  There is NO CPU fallback.
 */
#ifndef CKE_B200_H
#define CKE_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef int Int;     /* cke.hpp:4 */
typedef double Real; /* cke.hpp:5 */

/* ---- the reference interface, cke.hpp:7-23 ------------------------------------------------ */

/* cke.hpp:8 / cke.cpp:6-9: the reference starts Kokkos; here it selects the CUDA device
 * (LOCAL_RANK if set, else the current one) and creates the context. */
void kokkos_init(void);
/* cke.hpp:9 / cke.cpp:11-14: nothing to tear down beyond cke_cleanup. */
void kokkos_finalize(void);

/* cke.hpp:11-17 / cke.cpp:91-122,130-143: copies every input to the device (H2D inside, as the
 * reference's deep_copy), allocates a zeroed highOrderFlx(nvldim,nEdges) there.  nvldim >=
 * nVertLevels is the padded level extent (nested_vars.F90:107-108).  The host highOrderFlx
 * argument is not touched before cke_get_results. */
void cke_init(const Int nIters, const Int nEdges, const Int nCells, const Int nVertLevels, const Int nvldim,
              const Int nAdv, const Int *nAdvCellsForEdge, const Int *minLevelCell, const Int *maxLevelCell,
              const Int *advCellsForEdge, Real *tracerCur, const Real *normalThicknessFlux,
              const Real *advMaskHighOrder, const Real *cellMask, const Real *advCoefs, const Real *advCoefs3rd,
              const Real coef3rdOrder, Real *highOrderFlx);
/* cke.hpp:18 / cke.cpp:145-160: copies the device highOrderFlx (nvldim x nEdges of the cke_init
 * call; both size arguments are ignored exactly as the reference ignores them) to the host and
 * zeroes the device copy. */
void cke_get_results(const Int nEdges, const Int nVertLevels, Real *highOrderFlx);
/* cke.hpp:19 / cke.cpp:161 */
void cke_cleanup(void);

/* cke.hpp:21 / cke_impl1.cpp:11-47: nIters times { tracerCur *= cellMask in place;
 * highOrderFlx(k,e) = sum_i tracerCur(k,cell_i) * coef2(k,e) * (coef1_i + coef3_i*csgn(k,e)) },
 * then a fence. */
void cke_impl1_run(void);
/* cke.hpp:22 / cke_impl2.cpp:27-80 at pack size 1: nIters times the per-edge form in which each
 * contributing cell adds tracer*cellMask*wgt*(coef1+coef3*sgn) on its own level range only. */
void cke_impl2_run(void);

/* ---- the reference's YAKL-side interface, nested_loops/nested_yakl_c.cpp:130-306 ----------- */

/* nested_yakl_c.cpp:130-140: the input arrays, published by the Fortran driver as type(c_ptr), bind(C) globals
 * (nested_vars.F90:7, 75-86, 138-151).  Set them before yakl_init_arrays. */
extern int *c_maxLevelCell, *c_minLevelCell, *c_nAdvCellsForEdge, *c_advCellsForEdge;
extern double *c_advCoefs, *c_advCoefs3rd, *c_cellMask, *c_advMaskHighOrder, *c_tracerCur, *c_normalThicknessFlux, *c_highOrderFlx;
/* nested_yakl_c.cpp:142-146 */
void yakl_init(void);
/* nested_yakl_c.cpp:148-181: copies the connectivity, coefficients and masks behind the c_* pointers to the device */
void yakl_init_arrays(int fnvldim, int fnEdges, int fnCells, int fnVertLevels, int fnAdv, double fcoef3rdOrder);
/* nested_yakl_c.cpp:184-243: the original loop form (per-cell level ranges); uploads the host tracerCur(nvldim,nCells) and
 * normalThicknessFlux(nvldim,nEdges), computes, downloads highOrderFlx(nvldim,nEdges) -- all inside the call, as the reference */
void yakl_original_loop(double *h_highOrderFlx, double *h_tracerCur, double *h_normalThicknessFlux);
/* nested_yakl_c.cpp:260-306: the collapsed (edge, level) form with the cellMask factor */
void yakl_gpu_optimized(double *h_highOrderFlx, double *h_tracerCur, double *h_normalThicknessFlux);

/* ---- additions of this library (not in cke.hpp) -------------------------------------------- */

/* 1 (default): failures abort like the reference's asserts; 0: they return and set the message. */
void cke_b200_set_abort_on_error(int on);
/* NULL when the last cke_init and everything since succeeded. */
const char *cke_b200_last_error(void);
/* kernels launched by this library since cke_init */
long long cke_b200_launch_count(void);
/* Runs `iters` iterations of implementation `impl` (1 or 2) on the library stream between two
 * CUDA events and returns the mean milliseconds per iteration in *ms_per_iter; 0 on success. */
int cke_b200_time_ms(int impl, int iters, double *ms_per_iter);
/* Same, one iteration at a time with the L2 flushed before each (a 256 MB memset on the same
 * stream); returns the median of `reps` single-iteration times in *median_ms; 0 on success. */
int cke_b200_time_cold_ms(int impl, int reps, double *median_ms);
/* Copies the device tracerCur(nvldim,nCells) back (implementation 1 masks it in place). */
int cke_b200_get_tracer(Real *tracerCur);

#ifdef __cplusplus
}
#endif
#endif
