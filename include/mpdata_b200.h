/*
 * mpdata_b200.h -- C ABI of libmpdata_b200.so: the B200 (sm_100a) implementation of the MMF MPDATA
 * tracer step of the reference mini-app mmf-mpdata-tracer/ (SURVEY 8(f)4).
 *
 * Reference interface (mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90, cited as A:<line>):
 * three interchangeable internal subroutines with one argument list,
 *     advect_scalar2D_cpu / _openacc_1 / _openacc_2 (f, u, w, rho, rhow, flux)      A:477, A:247, A:72
 * plus the host-associated `adz` and the compile-time sizes nslices, nz, nx (A:6-8, A:29).  The
 * OpenACC forms create the device arrays, `update device` the inputs, run, and `update host(flux,f)`
 * (A:105-107, A:241).  mpdata_advect_scalar2D below is that whole region in one call; the
 * life-cycle calls split it for device-resident use.  Fortran binding: INTEGRATION.md section 7.
 *
 * Array layouts are the Fortran caller's (column-major, slice index fastest; A:14-36), nzm = nz-1:
 *   f   (nslices, -2:nx+3, 1, nzm)  inout     u (nslices, -1:nx+3, 1, nzm)  in     w (nslices, -1:nx+2, 1, nz)  in
 *   rho (nslices, nzm) in    rhow (nslices, nz) in    adz (nslices, nzm) in    flux (nslices, nz) out
 * flux(:, nz) is never written by the reference (A:477-642) and is not written here.
 *
 * Every function returns 0 on success; nonzero = failure, message from mpdata_last_error().
 * Failures: no sm_100 device (1), bad sizes (2; nz >= 3, nx >= 1, nslices >= 1), seven levels of a slice's
 * seven planes do not fit the 227 KB of shared memory of an SM (3; nx beyond ~570), CUDA runtime error (4), call order (5).
 * There is NO CPU fallback.  One calling thread; not re-entrant.
 */
#ifndef MPDATA_B200_H
#define MPDATA_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* A:72-244 as one call: allocate if the sizes changed, H2D of f,u,w,rho,rhow,adz, one kernel
 * launch, D2H of f and flux(:,1:nzm) into the caller's arrays. */
int mpdata_advect_scalar2D(int nslices, int nx, int nz, double *f, const double *u, const double *w, const double *rho,
                           const double *rhow, const double *adz, double *flux);

/* `!$acc enter data pcreate(...)` (A:105): device arrays for these sizes. */
int mpdata_init(int nslices, int nx, int nz);
/* `!$acc update device(f,u,w,rho,adz,rhow)` (A:107).  flux may be NULL (then the device copy of
 * level nz is zero). */
int mpdata_upload(const double *f, const double *u, const double *w, const double *rho, const double *rhow,
                  const double *adz, const double *flux);
/* One MPDATA step on the device arrays (f advances in place), asynchronous on the library stream. */
int mpdata_run(void);
/* `!$acc update host(flux,f)` (A:241); synchronises. */
int mpdata_download(double *f, double *flux);
int mpdata_finalize(void);

/* Measurement: `iters` steps, each from the uploaded f (restored device-to-device outside the
 * timed part), CUDA events around each launch on the library stream; flush_l2 != 0 writes a
 * buffer larger than L2 before each step.  *ms_per_step = mean kernel time. */
int mpdata_time_steps(int iters, int flush_l2, float *ms_per_step);
/* threads per CTA (multiple of 32, 32..1024; 0 = automatic, the default) -- results do not depend on it */
int mpdata_set_block(int threads);
/* level ranges a slice is split into (each CTA recomputes a halo of 3 levels either side); 0 = automatic, the
 * default: floor(SMs / nslices) while there are fewer slices than SMs, 1 otherwise, more when a whole slice does not
 * fit shared memory -- results do not depend on it */
int mpdata_set_chunks(int chunks);
/* Which kernel runs the step (results do not depend on it): 0 = chosen by the library, 1 = a CTA per slice (or level range)
 * with the seven phases separated by barriers, 2 = a warp per slice marching through the levels without any CTA-wide
 * barrier, 3 = the same with two slices per warp (2 and 3 are compiled for the reference sizes nx=32, nz=58 only:
 * mpdata_init fails for other sizes while one of them is selected).
 * warps_per_cta: for forms 2 and 3, 0 = the library's choice, else 1..12 (form 3: at most 6 are used). */
int mpdata_set_form(int form, int warps_per_cta);
/* The automatic choice (form 0) at the reference plane sizes: 1 if the marching kernel (form 2, 6 warps per CTA) is expected
 * to be faster than the barrier-phased one at `nslices` slices on a GPU of `sms` SMs, from the two kernels' measured
 * behaviour on a B200 (csrc/mpdata_b200.cu).  Host only, no GPU needed; MPDATA_B200_AUTO_MARCH=0 switches the choice off. */
int mpdata_march_wins(long long nslices, int sms);
/* the launch shape in use after mpdata_init (any pointer may be NULL) */
int mpdata_get_launch(int *grid, int *block, int *chunks, long long *smem_bytes);
/* bytes of shared memory a whole slice (chunks = 1) needs at these sizes */
long long mpdata_slice_smem_bytes(int nx, int nz);
/* kernels launched by this library since it was loaded */
long long mpdata_launch_count(void);
const char *mpdata_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
