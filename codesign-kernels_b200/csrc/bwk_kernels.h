// bwk_kernels.h -- launch interface between the C-ABI layer (bwk_api.cu) and the sm_100a kernels.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "bwk_internal.h"

namespace bwk {

constexpr int TILE_SLABS = 32;      // one warp, one slab per thread
constexpr int METRIC_DOUBLES = 64;  // folded metric: 4 doubles per GLL point
constexpr int ELEM_DOUBLES = 144;   // element_t, F90:23-27

enum SweepMode : int {
    MODE_LAPLACE = 0, // qtens -> qtens in place              (reference-pinned, F90:186-200)
    MODE_SWEEP1 = 1,  // qtens -> tmp                          (first laplace; neighbours read their edges straight from tmp)
    MODE_SWEEP2 = 2,  // tmp + neighbours' edge values -> *rspheremp -> laplace -> qtens  (edgeVunpack + second laplace)
    MODE_SWEEP1_PRE = 3, // as MODE_SWEEP1 on qtens * prescale(:,:,k,ie)  (advance_hypervis_scalar: Qtens = dpdiss*Qdp/dp)
    MODE_DSS = 4,     // as MODE_SWEEP2 without the Laplacian: tmp + neighbours' edge values -> out  (edgeVunpack [+ rspheremp])
    MODE_SWEEP2_POST = 5 // MODE_SWEEP2 with advance_hypervis_scalar's update fused in: out = limiter(Qdp*spheremp - coef*lap)*rspheremp
};

// Everything a sweep launch needs.  Tensor maps are passed by value (__grid_constant__).
struct SweepParams {
    CUtensorMap map_q;      // qtens   as {16, KQ, nelem}, box {16,32,1}, SWIZZLE_128B
    CUtensorMap map_tmp;    // tmp     same shape (sweep-1 output)
    CUtensorMap map_rows;   // tmp as {16, KQ, nelem}, box {4,32,1}, SWIZZLE_32B: one edge row (32 B) of 32 consecutive slabs
    CUtensorMap map_halo;   // received edge rows as {4, KQ, nrows}, box {4,32,1}, SWIZZLE_32B (aliases map_rows when there is no halo)
    CUtensorMap map_pt;     // tmp as {4 (i), 4 (j), KQ, nelem}, box {2,1,32,1}: the 16-byte i-pair holding one point of 32 consecutive slabs
    CUtensorMap map_halo_pt;// received edge rows as {4, KQ, nrows}, box {2,32,1}: half a halo row (aliases a 3-D view of tmp when there is no halo)
    CUtensorMap map_out;    // where sweep 2 / the DSS-only sweep store their result (qtens, or the hyperviscosity work buffer)
    // warp-strip sweep: boxes of 32/NM slabs x NM consecutive elements (NM = 4, 8 or 16 members per warp; set per launch)
    CUtensorMap map4_tmp;   // source field as {16, KQ, nelem}, box {16, 32/NM, NM}, SWIZZLE_128B
    CUtensorMap map4_rows;  // source field, box {4, 32/NM, NM}, SWIZZLE_32B: one edge row of the warp tile's slabs
    CUtensorMap map4_halo;  // received edge rows {4, KQ, nrows}, box {4, 32/NM, NM}, SWIZZLE_32B
    CUtensorMap map4_out;   // destination, box {16, 32/NM, NM}, SWIZZLE_128B
    CUtensorMap map4_q;     // qtens, same box (SK_POST reads Qdp)
    const double *prescale; // MODE_SWEEP1_PRE: [nelem][nlev][16] factors applied to the input slab of level k
    int nlev;               // MODE_SWEEP1_PRE: slab kq belongs to level kq % nlev
    const double *elem;     // MODE_SWEEP2_POST: element records [nelem][144] (spheremp at +64); Qdp is read through map_q
    double hv_coef;         // MODE_SWEEP2_POST: dt * nu_q
    int hv_limiter;         // MODE_SWEEP2_POST: apply limiter2d_zero
    const double *metric;   // [nelem][64] folded metric
    const double *rsph;     // [nelem][16]
    const ElemDss *dss;     // [nelem]
    const double *tmp;      // sweep-1 output, natural layout (neighbours' edge rows/columns are read from here)
    const double *halo;     // received edge rows [nrows][KQ][4]
    const int32_t *elem_list; // optional list of local element indices, else contiguous from elem_first
    const int32_t *strips;  // MODE_SWEEP2: work units, first*8 + (count-1) (bwk_internal.h); grid x = strip_count
    const StripReg *strip_reg; // MODE_SWEEP2: one per strip, same order
    int strip_count;
    int *work_counter;      // persistent strip sweep: global item counter, zero at launch
    int debug_flags;        // warp-strip sweep: timing experiments (BWK_DEBUG_FLAGS), 0 in production
    int nbr_ahead;          // persistent strip sweep: the S/N rows are requested this many tiles ahead (1 or 2)
    int elem_first;         // 0-based first local element (when elem_list == nullptr)
    int elem_count;         // grid x (MODE_LAPLACE / MODE_SWEEP1)
    int kq;                 // nlev*qsize
    int tiles_per_unit;     // tiles (of 32 slabs) per CTA work unit
    int units;              // grid y
};

// geometry chosen by the host for a given KQ / element count
struct SweepGeometry {
    int tiles_per_unit, units;
};
// tpu_hint: tiles per unit wanted for the strip sweeps (0 = default 15)
SweepGeometry choose_geometry(int kq, int64_t nelem, int mode, int tpu_hint = 0);
void set_sweep2_tiles_per_unit(int tpu); // tuning knob (BWK_SWEEP2_TPU): > 0 forces the unit length and disables the autotuner
int sweep2_tiles_per_unit_forced();

cudaError_t launch_sweep(int mode, const SweepParams &p, cudaStream_t st);
// warp-strip sweep (MODE_SWEEP2 / MODE_DSS / MODE_SWEEP2_POST): grid (p.strip_count, p.units), p.tiles_per_unit tiles of 32 slabs per CTA
cudaError_t launch_sweep2_wstrip(int mode, const SweepParams &p, cudaStream_t st, int members = 4);
// persistent strip sweep (MODE_SWEEP2 / MODE_DSS) over full regular strips only: p.strip_reg / p.strip_count, p.units units of >= 3 tiles
cudaError_t launch_sweep2_persist(int mode, const SweepParams &p, cudaStream_t st);
// MODE_SWEEP1_PRE needs nlev*128 more bytes of shared memory per CTA than the other element-local sweeps; false if that does not fit
bool sweep1_pre_fits(int nlev);
// advance_hypervis_scalar, pointwise part: out = limiter2d_zero(qdp*spheremp - coef*qtens) * rspheremp over the slabs of
// local elements [first, first+count) or of a device list (HOMME prim_advection, not in the reference)
cudaError_t launch_hypervis_post(const double *qdp, const double *qtens, double *out, const double *elem, const double *rsph,
                                 const int32_t *list, int first, int count, int kq, double coef, int limiter, cudaStream_t st);
// FP64 FMA issue-rate microbenchmark: blocks x 256 threads x 8 chains x 4096 FMAs
cudaError_t launch_dfma_peak(double *out, int blocks, cudaStream_t st);
// prescale = dpdiss / dp, n values
cudaError_t launch_ratio(const double *dpdiss, const double *dp, double *ratio, int64_t n, cudaStream_t st);
// plain-load variant of MODE_LAPLACE (no TMA, no shared-memory staging); for ncu A/B only
cudaError_t launch_laplace_plain(double *qtens, const double *metric, int elem_first, int elem_count, int kq, cudaStream_t st);

cudaError_t set_dvv(const double *dvv16_host, cudaStream_t st);
// folded metric M = -(rrearth^2) * spheremp * Dinv * tensorVisc * Dinv^T per GLL point
cudaError_t launch_fold_metric(const double *elem, double *metric, int64_t nelem, cudaStream_t st, int use_tensor = 1);
// element-local part of biharmonic_wk_dp3d over [elem_first, elem_first+elem_count) or a device list (see bwk_kernels.cu)
cudaError_t launch_dp3d_laplace(bool produce, const double *src, double *dst, const double *metric, const double *v2c, const double *elem,
                                const double *rsph, const int32_t *list, int elem_first, int elem_count, int nlev, int undamp, cudaStream_t st);
// edge pack: copy rows (local element, row id) x KQ into a contiguous send buffer (HOMME edgeVpack analogue)
cudaError_t launch_edge_pack(const double *tmp, const int32_t *row_elem, const int32_t *row_id,
                             int64_t nrows, int kq, double *sendbuf, cudaStream_t st);
// harness: initialize_data on the device (F90:48-58), table = the 131072 LCG draws
cudaError_t launch_init_data(const double *table, int64_t nelemd_global, int64_t elem_first, int64_t nelem, int kq,
                             int seed_mode, double *elem, double *qtens, cudaStream_t st);
cudaError_t launch_l2_partials(const double *q, const double *qs, int64_t n, double *partials /*[2*nblocks]*/, int nblocks, cudaStream_t st);
cudaError_t launch_sum_partials(const double *q, int64_t n, double *partials /*[2*nblocks]*/, int nblocks, cudaStream_t st);
// bit checksum: partials[b] = sum over the block's values of mix64(bits) * (2*(index0 + t) + 1) mod 2^64
cudaError_t launch_bits_checksum(const double *q, int64_t n, uint64_t index0, unsigned long long *partials /*[nblocks]*/, int nblocks, cudaStream_t st);
cudaError_t launch_period_defect(const double *q, int64_t n, int64_t period, double *partials /*[nblocks]*/, int nblocks, cudaStream_t st);

} // namespace bwk
