/*
 * bwk_b200.h -- C ABI of libbwk_b200.so: the B200 (sm_100a) implementation of the HOMME
 * scalar hyperviscosity operator extracted in the reference file
 *     atmosphere/biharmonic_wk_kernel.F90                    (cited as F90:<line>)
 *
 * This header is the drop-in boundary.  Everything is extern "C", plain pointers and
 * sizes; arrays keep the reference's Fortran column-major layouts bit for bit, so a
 * Fortran caller passes its arrays unchanged (see fortran/biharmonic_wk_scalar_b200.F90
 * and INTEGRATION.md for the ISO_C_BINDING interface block).
 *
 * Conventions
 *   - every function returns 0 on success, nonzero on failure; bwk_last_error() then
 *     returns a message.  There is NO CPU fallback: without a usable CUDA device every
 *     compute entry point fails.
 *   - one calling thread, one context per process (the reference keeps its arrays as
 *     module-level variables, F90:29-33); with several GPUs use one process per GPU.
 *   - nets/nete are 1-based inclusive element indices into the uploaded elements,
 *     exactly as in F90:186-191.
 */
#ifndef BWK_B200_H
#define BWK_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BWK_NP 4

/* type element_t, F90:23-27.  C index order is the reverse of Fortran's:
 * Dinv(i,j,c,a) == Dinv[a-1][c-1][j-1][i-1].  144 doubles, no padding. */
typedef struct bwk_element {
    double Dinv[2][2][BWK_NP][BWK_NP];
    double spheremp[BWK_NP][BWK_NP];
    double tensorVisc[2][2][BWK_NP][BWK_NP];
} bwk_element_t;

/* type derivative_t, F90:19-21.  Dvv(i,l) == Dvv[l-1][i-1]. */
typedef struct bwk_derivative {
    double Dvv[BWK_NP][BWK_NP];
} bwk_derivative_t;

/* ------------------------------------------------------------------------------------
 * Reference-pinned path: one weak Laplacian with tensor viscosity, in place (F90:186-200).
 * ---------------------------------------------------------------------------------- */

/* `allocate_data` + `!$acc enter data pcreate(qtens,elem,deriv,grads)`, F90:37-44,553.
 * Creates the device mirrors for nelemd elements of (np,np,nlev,qsize).  np is 4;
 * nlev, qsize, nelemd are run-time here (compile-time parameters in F90:11-17). */
int bwk_init(int nelemd, int nlev, int qsize);
/* releases every device/pinned-host resource of the context */
int bwk_finalize(void);

/* `!$acc update device(qtens,elem,deriv)`, F90:573.  Any pointer may be NULL to skip it.
 * Uploading elem+deriv also rebuilds the folded per-point metric on the device. */
int bwk_upload(const bwk_element_t *elem, const double *qtens, const bwk_derivative_t *deriv);
/* `call bws_gpu2(elem,qtens,grads,deriv,nets,nete)`, F90:574 / F90:525-534, on the device
 * mirrors; asynchronous on the library stream. */
int bwk_run(int nets, int nete);
/* `!$acc update host(qtens)`, F90:575 (synchronises). */
int bwk_download(double *qtens);
int bwk_sync(void);

/* The reference's whole timed region F90:572-576 in one call, GPU2 signature F90:525-531:
 *   subroutine biharmonic_wk_scalar(elem,qtens,grads,deriv,nets,nete)
 * elem(nelemd), qtens(np,np,nlev,qsize,nelemd), grads ignored (may be NULL), scalars by
 * reference as Fortran passes them.  Upload, compute and download are pipelined over
 * element chunks.  Requires bwk_init with the same nelemd/nlev/qsize. */
int bwk_biharmonic_wk_scalar(const bwk_element_t *elem, double *qtens, double *grads,
                             const bwk_derivative_t *deriv, const int *nets, const int *nete);
/* CPU/GPU1 signature F90:186-191 / F90:317-323: qtens(np,np,nlev,qsize,nets:nete), i.e. the
 * first slab passed is element `nets`; elem is still the whole elem(:) array. */
int bwk_biharmonic_wk_scalar_nets(const bwk_element_t *elem, double *qtens_nets,
                                  const bwk_derivative_t *deriv, const int *nets, const int *nete);

/* Pinned host memory for the caller's arrays (the reference builds with -ta=...,pinned,
 * atmosphere/Makefile:17-20).  Optional: pageable memory works, slower. */
int bwk_host_alloc(void **ptr, uint64_t bytes);
int bwk_host_free(void *ptr);
int bwk_host_register(void *ptr, uint64_t bytes);
int bwk_host_unregister(void *ptr);

/* Allocation-free form of the single Laplacian above, for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * Needs no bwk_init and touches no context state.  d_elem (nelemd element_t records) and d_qtens
 * (np,np,nlev,qsize,nelemd; 128-byte aligned; updated in place as F90:186-200) are DEVICE pointers of the current device,
 * deriv is a HOST pointer (64 bytes, staged into the library's __constant__ Dvv as bwk_upload does), stream is a
 * cudaStream_t as an integer (0 = the default stream); the two kernels are queued on it and the call returns without
 * synchronising.  The folded per-point metric, which bwk_init keeps in a cudaMalloc'd buffer, lives here in a file-scope
 * static __device__ array of BWK_NOALLOC_MAX_NELEMD * 64 doubles (4 MiB, held by every GPU that loads the library until the
 * process exits); nelemd beyond that is rejected with an error.  The call makes no cudaMalloc / cudaFree and calls no
 * allocating library routine.  All calls share the one array and each rewrites the part it reads, so nothing has to be
 * cleared between calls, but two calls that overlap on the GPU (different streams, or a graph replayed concurrently) race
 * on it, and so does a concurrent bwk_upload of another Dvv: serialise them. */
#define BWK_NOALLOC_MAX_NELEMD 8192
int bwk_biharmonic_wk_scalar_noalloc(const bwk_element_t *d_elem, double *d_qtens, const bwk_derivative_t *deriv,
                                     int nelemd, int nlev, int qsize, uint64_t stream);

const char *bwk_last_error(void);
/* kernel variant for bwk_run: 0 = TMA-pipelined (default), 1 = plain loads (kept for ncu A/B) */
int bwk_set_variant(int variant);
/* number of kernels this library has launched since bwk_init (for bench.py gpu_launches) */
int64_t bwk_launch_count(void);
/* the cudaStream_t all kernels are launched on, as an integer handle */
int bwk_get_stream(uint64_t *stream);
/* device pointer of the qtens mirror (for zero-copy wrappers in tests/bench) */
int bwk_get_device_qtens(uint64_t *ptr);
/* CUDA-event timing on the library stream (the analogue of the system_clock pairs F90:556-558,
 * 572-576, but on the device): 64 slots; elapsed blocks until event `b` has happened. */
int bwk_event_record(int slot);
int bwk_event_elapsed_ms(int a, int b, float *ms);
/* Single rank: record event `slot` on the library stream between the element-local sweep and the strip sweep of every
 * DSS-containing call that follows (slot < 0 switches it off), so that a harness can time the two sweeps INSIDE its timed
 * steps: bwk_event_elapsed_ms(start, slot) and (slot, end). */
int bwk_set_phase_event(int slot);

/* ------------------------------------------------------------------------------------
 * Harness helpers, mirroring module kernel_utils (F90:7-93) on the device so that the
 * large configurations never need a host copy.
 * ---------------------------------------------------------------------------------- */

/* `initialize_data`, F90:48-58 + myrandom F90:77-91, generated on the device.  With
 * several ranks the values are those of the GLOBAL arrays: nelemd_global elements in
 * total, of which this context holds [elem_first, elem_first+nelemd).  seed_mode 0 = the
 * reference LCG (period 131072); 1 = splitmix64 high-entropy values in [0,1) (robustness
 * runs only, not reference data). */
int bwk_initialize_data_device(int64_t nelemd_global, int64_t elem_first, int seed_mode);
/* `save_qtens`, F90:62-65 (device-side copy) */
int bwk_save_qtens(void);
/* `compute_l2norm`, F90:69-73, of the device qtens against the saved copy */
int bwk_compute_l2norm(double *l2);
/* sum and sum of squares of the device qtens (checksums for size-independent tests) */
int bwk_checksum(double *sum, double *sumsq);
/* checksum of the BITS of the device qtens that does not depend on how the elements are split over ranks:
 * sum over values of mix64(bits) * (2*global_index + 1) mod 2^64, global_index counted from global element
 * `elem_first_global`.  The per-rank values of any contiguous partition add up (mod 2^64) to the 1-rank value,
 * so equal sums prove "N ranks == 1 rank bit for bit" (bench.py prints it for every N). */
int bwk_checksum_bits(int64_t elem_first_global, uint64_t *hash);
/* max_i |q_i - q_{i mod period}| over the device qtens, period in elements (LCG inputs
 * repeat every 8192 elements, so must the reference-pinned output) */
int bwk_periodicity_defect(int64_t period_elems, double *max_abs_diff);
/* copy nelem elements starting at 0-based element `first` of the device qtens to host */
int bwk_download_elements(int64_t first, int64_t count, double *dst);
int bwk_upload_elements(int64_t first, int64_t count, const double *src);

/* ------------------------------------------------------------------------------------
 * Full biharmonic of upstream HOMME (viscosity_base.F90 biharmonic_wk_scalar; only the
 * comment F90:518-519 survives in the reference -> parity unpinned):
 *   qtens <- laplace_sphere_wk(qtens); edgeVpack; bndry_exchangeV; edgeVunpack;
 *   qtens <- laplace_sphere_wk(rspheremp * qtens)
 * ---------------------------------------------------------------------------------- */

/* Connectivity as global degree-of-freedom ids, one per GLL point of every GLOBAL element:
 * gdof[16*e + i + 4*j].  Points with equal id are the same node (HOMME: elem%gdofP).  The
 * context's elements are global elements [elem_first, elem_first+nelemd).  Builds the DSS
 * plan (edge/corner sources, halo send/recv lists) and rspheremp. */
int bwk_set_topology_gdof(int64_t nelem_global, const int64_t *gdof, int64_t elem_first);
/* The synthetic cubed sphere of SURVEY Appendix D: nelem_global = 6*ne*ne, element id
 * f*ne*ne + je*ne + ie. */
int bwk_set_topology_cubed_sphere(int ne, int64_t elem_first);
/* fills gdof (16*6*ne*ne entries) for that cubed sphere; host only, no GPU needed */
int bwk_cubed_sphere_gdof(int ne, int64_t *gdof);
/* optional: caller-supplied rspheremp(np,np,nelemd) instead of the DSS-derived one */
int bwk_set_rspheremp(const double *rspheremp);
int bwk_get_rspheremp(double *rspheremp);

/* Laplace -> DSS -> *rspheremp -> Laplace on all local elements; asynchronous. */
int bwk_run_biharmonic(void);
/* one phase of the above on a single rank, for per-kernel timing and profiling only:
 * 1 = first sweep (laplace into the scratch field), 2 = second sweep (unpack, rspheremp, laplace) */
int bwk_run_biharmonic_phase(int phase);
/* same through host buffers (upload, both sweeps, download), GPU2-style signature */
int bwk_biharmonic_wk_scalar_full(const bwk_element_t *elem, double *qtens,
                                  const bwk_derivative_t *deriv);

/* ---- the caller of this path in HOMME: advance_hypervis_scalar ----------------
 * NOT in the reference repository (upstream E3SM components/homme/src/share/prim_advection_base.F90, un-vendored, no
 * pinned version; semantics restated in oracle/bwk_oracle.c, PARITY UNPINNED).  It replaces, per subcycle
 * (dt = dt2 / hypervis_subcycle_q):
 *     Qtens = dpdiss * Qdp / dp ; call biharmonic_wk_scalar(elem, Qtens, deriv, edge, hybrid, nets, nete)
 *     Qdp   = Qdp * spheremp - dt * nu_q * Qtens ; limiter2d_zero(Qdp(:,:,:,q)) ; edgeVpack / bndry_exchangeV /
 *     edgeVunpack ; Qdp = Qdp * rspheremp
 * Qdp(np,np,nlev,qsize,nelemd) is the field given to bwk_upload / held in the device buffer; dp and dpdiss are
 * (np,np,nlev,nelemd): dp = derived%dp - dt2*derived%divdp_proj, dpdiss = derived%dpdiss_ave (or dp0(k) broadcast).
 * Needs a topology (bwk_set_topology_*).  Allocates a second field-sized work buffer on first use. */
int bwk_upload_hypervis_fields(const double *dp, const double *dpdiss);
/* device resident, asynchronous; limiter != 0 applies limiter2d_zero */
int bwk_run_advance_hypervis(double dt2, double nu_q, int hypervis_subcycle_q, int limiter);
/* upload (elem, Qdp, dp, dpdiss, deriv), run, download Qdp */
int bwk_advance_hypervis_scalar(const bwk_element_t *elem, double *Qdp, const double *dp, const double *dpdiss,
                                const bwk_derivative_t *deriv, double dt2, double nu_q, int hypervis_subcycle_q, int limiter);

/* ---- biharmonic_wk_dp3d: the momentum / temperature / layer-thickness counterpart ----
 * NOT in the reference repository (upstream HOMME viscosity_mod: biharmonic_wk_dp3d(elem, dptens, dpflux, ttens, vtens,
 * deriv, edge3, hybrid, nt, nets, nete); derivative_mod: vlaplace_sphere_wk / vlaplace_sphere_wk_cartesian; un-vendored,
 * no pinned version; semantics restated in oracle/bwk_oracle.c, PARITY UNPINNED, pinned to analytic eigenvalues by
 * tests/test_dp3d.py).  It replaces
 *     ttens = laplace_sphere_wk(T) ; dptens = laplace_sphere_wk(dp3d) ; vtens = vlaplace_sphere_wk(v)    [var_coef1]
 *     edgeVpack(ttens @0, vtens @nlev, dptens @3 nlev) / bndry_exchangeV / edgeVunpack ; * rspheremp
 *     the same three operators again                                                                      [var_coef = .true.]
 * with vlaplace_sphere_wk in its tensor-hyperviscosity (Cartesian) form: lat-lon components -> three Cartesian components
 * through vec_sphere2cart(np,np,3,2), laplace_sphere_wk (F90:164-182) on each, back with the transpose, and, when
 * undamprrcart != 0, + 2*spheremp*v*rrearth**2 so that rigid rotation is not damped.  var_coef1 = 0 leaves tensorVisc out
 * of the first application (upstream does that when hypervis_scaling > 0); the contravariant (grad-div / curl-vorticity)
 * form and the dpflux output are not provided.
 * The context must be made with bwk_init(nelemd, nlev, 4): the device field holds, per element, T(np,np,nlev),
 * v(np,np,2,nlev), dp3d(np,np,nlev) in that order (HOMME's edge-buffer order).  On entry ttens / dptens / vtens hold T /
 * dp3d / v (as qtens holds the input of biharmonic_wk_scalar), on return the weak biharmonic. */
int bwk_upload_vec_sphere2cart(const double *vec_sphere2cart /* (np,np,3,2,nelemd) */);
/* device resident (the fused field is the one bwk_upload / bwk_download move), asynchronous */
int bwk_run_biharmonic_dp3d(int var_coef1, int undamprrcart);
/* upload (elem, vec_sphere2cart, fields, deriv), run, download; ttens, dptens (np,np,nlev,nelemd), vtens (np,np,2,nlev,nelemd);
 * elem / vec_sphere2cart / deriv may be NULL to keep what is on the device */
int bwk_biharmonic_wk_dp3d(const bwk_element_t *elem, const double *vec_sphere2cart, double *dptens, double *ttens, double *vtens,
                           const bwk_derivative_t *deriv, int var_coef1, int undamprrcart);

/* ---- several GPUs: one process per GPU, contiguous element blocks, NCCL send/recv ---- */
#define BWK_UNIQUE_ID_BYTES 128
/* rank 0 creates the id, the launcher broadcasts the bytes (any transport) */
int bwk_comm_unique_id(void *id_bytes);
int bwk_comm_init(int rank, int nranks, const void *id_bytes);
/* harness: measured FP64 FMA rate of this GPU in TFLOP/s (DFMA microbenchmark), the denominator for FP64 pipe utilisation */
int bwk_measure_fp64_peak(double *tflops);
/* device-side barrier across ranks on the library's stream (benchmark bracketing); no-op on one rank */
int bwk_comm_barrier(void);
int bwk_comm_finalize(void);
/* contiguous block partition used by every rank: [first, first+count) */
int bwk_partition(int64_t nelem_global, int rank, int nranks, int64_t *first, int64_t *count);

/* ---- DSS plan introspection (host only; used by the CPU tests, no GPU needed) ---- */
typedef struct bwk_plan bwk_plan_t;
bwk_plan_t *bwk_plan_create(int64_t nelem_global, const int64_t *gdof, int rank, int nranks);
void bwk_plan_destroy(bwk_plan_t *);
int64_t bwk_plan_unique_nodes(const bwk_plan_t *);
int64_t bwk_plan_local_first(const bwk_plan_t *);
int64_t bwk_plan_local_count(const bwk_plan_t *);
/* number of halo rows received from / sent to `peer` (rows of 4 points x nlev*qsize) */
int64_t bwk_plan_recv_rows(const bwk_plan_t *, int peer);
int64_t bwk_plan_send_rows(const bwk_plan_t *, int peer);
/* send list: pairs (global element, row id) in wire order; row ids 0=S(j=1) 3=N(j=4) 4=W(i=1) 5=E(i=4) */
int bwk_plan_send_list(const bwk_plan_t *, int peer, int64_t *elem, int32_t *row);
int bwk_plan_recv_list(const bwk_plan_t *, int peer, int64_t *elem, int32_t *row);
int64_t bwk_plan_boundary_elems(const bwk_plan_t *);  /* local elements that feed a peer  */
int64_t bwk_plan_dependent_elems(const bwk_plan_t *); /* local elements that need a halo */
/* sweep-2 work units ("strips": up to 4 consecutive local elements chained W->E whose shared edges
 * are exchanged on chip).  which: 0 = all local elements, 1 = halo-dependent, 2 = halo-independent. */
int64_t bwk_plan_strip_count(const bwk_plan_t *, int which);
/* regular[i] = 1 if strip i takes the fully asynchronous structured-patch path (may be NULL) */
int bwk_plan_strip_list(const bwk_plan_t *, int which, int32_t *first_local_elem, int32_t *count, int32_t *regular);
/* How install_plan groups list `which` for the warp-strip sweep when a warp may carry up to max_members
 * elements (4, 8 or 16): returns the number of work units and, when first != NULL, writes for each the first local element and
 * the number of chained full strips in it (4, 2 or 1; 0 = a strip the warp-strip sweep does not take, which runs in the
 * round-1 strip kernel).  Both arrays need bwk_plan_strip_count(plan, which) entries.  Host only; used by the CPU tests. */
int64_t bwk_plan_strip_groups(const bwk_plan_t *, int which, int max_members, int32_t *first_local_elem, int32_t *nstrips);
/* CPU evaluation of the plan's gather for one slab field (16 doubles per GLOBAL element in
 * `field`): writes the DSS sums of the local elements to out (16*local_count).  Test hook
 * proving the plan equals a node-wise sum; not used by the GPU path. */
int bwk_plan_apply_host(const bwk_plan_t *, const double *field, double *out);

#ifdef __cplusplus
}
#endif
#endif
