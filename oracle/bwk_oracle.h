/*
 * bwk_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the hot path in the reference file
 *   atmosphere/biharmonic_wk_kernel.F90            (cited below as F90:<line>)
 * The reference is Fortran and no Fortran compiler exists in this image.  This
 * restatement is pinned twice: against the two known-answer scalars the reference
 * publishes in atmosphere/results/ (tests/golden/kat.json, tests/test_oracle_kat.py),
 * and BIT FOR BIT against the reference's own three modules translated mechanically
 * from the Fortran to C and compiled here (oracle/f90_to_c.py -> oracle/_ref/libatm_f90ref.so,
 * oracle/ref_atm_py.py, tests/test_zzy_atm_ref.py) -- a translation that itself
 * reproduces those published scalars to all 17 digits.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (libbwk_b200.so) never does.
 *
 * Parity status:
 *   - single weak Laplacian (F90:100-202 and the GPU1/GPU2 orders):  PINNED by KAT-1/KAT-2 and by the translated reference.
 *   - DSS + rspheremp + second Laplacian (upstream HOMME viscosity_base/edge_mod_base,
 *     not vendored under /root/reference, only the comment F90:518-519):  PARITY UNPINNED.
 *
 * Array layouts are the reference's Fortran column-major ones, flattened 0-based:
 *   qtens[i + 4*j + 16*k + 16*nlev*q + 16*nlev*qsize*ie]               F90:31,41
 *   elem[144*ie + ...] : Dinv(i,j,c,a) at  i+4j+16c+32a      (64)      F90:24
 *                        spheremp(i,j)  at 64 + i+4j          (16)      F90:25
 *                        tensorVisc(i,j,c,a) at 80 + i+4j+16c+32a (64)  F90:26
 *   Dvv(i,l) at i + 4*l                                                 F90:20
 */
#ifndef BWK_ORACLE_H
#define BWK_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BWKO_NP 4
#define BWKO_ELEM_DOUBLES 144

/* F90:14 -- the literal has no kind suffix, so it is rounded to real(4) first. */
double bwko_rrearth(void);

/* F90:77-91.  state is the SAVEd integer `old`; reset != 0 sets it to 11 first. */
void bwko_myrandom(int64_t n, double *a, int reset, int *state);

/* F90:48-58 with nets=1, nete=nelemd (F90:15-17): Dvv (reset), then per element
 * Dinv, spheremp, tensorVisc, then qtens. */
void bwko_initialize_data(int nelemd, int nlev, int qsize,
                          double *dvv, double *elem, double *qtens);

/* Same values as bwko_initialize_data(nelemd_total, ...) but only elements
 * [first, first+count) are materialised (elem and qtens hold `count` elements).
 * Uses the closed form: draw d (1-based) = table[(d-1) mod 131072]. */
void bwko_initialize_subset(int64_t nelemd_total, int nlev, int qsize,
                            int64_t first, int64_t count,
                            double *dvv, double *elem, double *qtens);

/* F90:186-200 (module biharmonic_wk_scalar_cpu), operation order preserved.
 * nets/nete are 1-based inclusive; qtens and elem are indexed from element 1. */
void bwko_biharmonic_wk_scalar_cpu(const double *elem, double *qtens, const double *dvv,
                                   int nlev, int qsize, int nets, int nete);
/* Same arithmetic, `#pragma omp parallel for` over ie (builder extension). */
void bwko_biharmonic_wk_scalar_cpu_omp(const double *elem, double *qtens, const double *dvv,
                                       int nlev, int qsize, int nets, int nete);
int bwko_omp_max_threads(void);

/* F90:317-360 (GPU1, compiler-inline) operation order; nlev must be a multiple of 8. */
void bwko_biharmonic_wk_scalar_gpu1(const double *elem, double *qtens, const double *dvv,
                                    int nlev, int qsize, int nets, int nete);
/* F90:525-534 (GPU2, push-loop) operation order; grads is (4,4,2,nlev,qsize,nelemd) scratch. */
void bwko_biharmonic_wk_scalar_gpu2(const double *elem, double *qtens, double *grads,
                                    const double *dvv, int nlev, int qsize, int nelemd,
                                    int nets, int nete);

/* F90:69-73  sqrt( sum((q-qs)^2) / sum(qs^2) ), sequential sums in array order. */
double bwko_compute_l2norm(const double *qtens, const double *qtens_save, int64_t n);
/* max_i |a_i-b_i| / |b_i|  (b_i == 0 -> |a_i|) */
double bwko_max_rel(const double *a, const double *b, int64_t n);

/* ---------- DSS + second sweep: upstream HOMME semantics, PARITY UNPINNED ---------- */

/* SURVEY Appendix D lattice cube: gdof[16*e + i + 4*j] = x*(L+1)^2 + y*(L+1) + z,
 * L = 3*ne, global element id e = f*ne*ne + je*ne + ie. nelem = 6*ne*ne. */
void bwko_cubed_sphere_gdof(int ne, int64_t *gdof);

typedef struct bwko_dss bwko_dss_t;
/* Groups the 16*nelem points by gdof id; members of a group are kept in ascending
 * (element, point) order, which fixes the summation order. */
bwko_dss_t *bwko_dss_create(int64_t nelem, const int64_t *gdof);
void bwko_dss_destroy(bwko_dss_t *);
int64_t bwko_dss_unique_nodes(const bwko_dss_t *);
/* histogram of node multiplicities, mult[m] = number of nodes shared by m points, m<16 */
void bwko_dss_multiplicity_hist(const bwko_dss_t *, int64_t *mult16);
/* rspheremp(e,pt) = 1 / sum_{sharing points} spheremp   (rsph has 16*nelem doubles) */
void bwko_dss_rspheremp(const bwko_dss_t *, const double *elem, double *rsph);
/* In-place direct stiffness summation of every (k,q) slab: each shared point
 * becomes the sum over its group, identical in every sharing element. */
void bwko_dss_apply(const bwko_dss_t *, double *qtens, int nlev, int qsize);

/* Laplace (CPU order) -> DSS -> *rspheremp -> Laplace (CPU order), all elements. */
void bwko_biharmonic_full(const bwko_dss_t *, const double *elem, const double *rsph,
                          double *qtens, const double *dvv, int nlev, int qsize,
                          int use_omp);

/* advance_hypervis_scalar of upstream HOMME (the caller of biharmonic_wk_scalar; NOT in the reference,
 * parity unpinned): per subcycle Qtens = dpdiss*Qdp/dp -> biharmonic -> Qdp*spheremp - dt*nu_q*Qtens ->
 * limiter2d_zero (optional) -> DSS -> *rspheremp.  qdp(np,np,nlev,qsize,nelemd); dp, dpdiss (np,np,nlev,nelemd). */
void bwko_advance_hypervis_scalar(const bwko_dss_t *, const double *elem, const double *rsph, double *qdp,
                                  const double *dp, const double *dpdiss, const double *dvv, int nlev, int qsize,
                                  double dt2, double nu_q, int subcycles, int limiter);
/* limiter2d_zero on nslabs (np x np) levels: negatives clipped, the level's sum preserved */
void bwko_limiter2d_zero(double *q, int64_t nslabs);

/* biharmonic_wk_dp3d / vlaplace_sphere_wk (Cartesian form) of upstream HOMME (NOT in the reference, parity unpinned; see
 * bwk_oracle.c).  ttens, dptens (np,np,nlev,nelemd) hold T and dp3d on entry, vtens (np,np,2,nlev,nelemd) holds v; all are
 * overwritten with the weak biharmonic.  v2c = vec_sphere2cart (np,np,3,2,nelemd). */
void bwko_biharmonic_wk_dp3d(const bwko_dss_t *, const double *elem, const double *v2c, const double *rsph, double *dptens,
                             double *ttens, double *vtens, const double *dvv, int nlev, int var_coef1, int undamp);
/* one vlaplace_sphere_wk (Cartesian form) per (level, element), in place on v (np,np,2,nlev,nelemd) */
void bwko_vlaplace_sphere_wk(const double *elem, const double *v2c, double *v, const double *dvv, int nlev, int64_t nelem,
                             int var_coef, int undamp);

#ifdef __cplusplus
}
#endif
#endif
