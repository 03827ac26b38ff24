// bwk_kernels.cu -- hand-written FP64 kernels for sm_100a (B200).
//
// The operator is the weak-form spherical Laplacian with tensor viscosity of the reference
// file atmosphere/biharmonic_wk_kernel.F90 (gradient_sphere F90:109-134, tensor step
// F90:172-180, divergence_sphere_wk F90:138-160), evaluated in the folded form of SURVEY
// Appendix A:
//     dx(i,j) = sum_l Dvv(l,i) s(l,j)          dy(i,j) = sum_l Dvv(l,j) s(i,l)
//     (w1,w2) = M(i,j) (dx,dy)                 M = -rrearth^2 spheremp Dinv tensorVisc Dinv^T
//     out(i,j) = sum_l [ w1(l,j) Dvv(i,l) + w2(i,l) Dvv(j,l) ]
// which is 20 FP64 FMA-class instructions per GLL point instead of the reference's 46.
//
// Design (B200-first, not a translation of the OpenACC loops):
//   * one thread owns one (np x np) slab = 128 contiguous bytes; all 4x4 contractions are
//     register-resident, no shuffles, no shared-memory reductions;
//   * a warp owns a tile of 32 consecutive slabs (4 KB) of one element and runs its own
//     TMA pipeline: cp.async.bulk.tensor loads into SWIZZLE_128B shared memory (so the
//     per-thread 128-byte reads are bank-conflict free), results are written back in place
//     in shared memory and leave through a TMA store -- no LSU global traffic at all;
//   * the per-element folded metric (64 doubles) is staged once per CTA in shared memory
//     and read as warp-wide broadcasts; Dvv lives in __constant__ memory and is consumed
//     directly as a constant-bank operand of DFMA;
//   * sweep 2 of the full biharmonic adds the neighbours' coincident edge values (DSS, HOMME
//     edgeVunpack) read with one 256-bit load per edge straight from the neighbours' sweep-1
//     output / W-E edge pack / halo buffer.
#include "bwk_kernels.h"

#include <algorithm>
#include <cstddef>
#include <cstdio>
#include <cstdlib>

namespace bwk {

// c_dvv[i + 4*l] == Dvv(i+1,l+1) of type derivative_t (F90:19-21)
__constant__ double c_dvv[16];

// the arithmetic (laplace_slab, fold_metric_point, BWK_RREARTH): shared with the host build of tests/driver/bwk_math_host.cpp
#include "bwk_math.cuh"

// ------------------------------------------------------------------------------------------
// PTX helpers (mbarrier + TMA)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) {}
}
__device__ __forceinline__ void fence_barrier_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, int c3, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *map, int c0, int c1, int c2, uint32_t src)
{
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3}], [%4];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(src)
                 : "memory");
}
// one arrival (release at CTA scope: the thread's earlier shared-memory writes are visible to whoever waits on the phase)
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void tma_store_3d_hint(const CUtensorMap *map, int c0, int c1, int c2, uint32_t src, uint64_t pol)
{
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group.L2::cache_hint [%0, {%1, %2, %3}], [%4], %5;"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(src), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait()
{
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

struct SmemMetric {
    const double *m;
    __device__ __forceinline__ double2 operator()(int pt, int half) const
    {
        return *reinterpret_cast<const double2 *>(m + 4 * pt + 2 * half);
    }
};
struct GlobalMetric {
    const double *m;
    __device__ __forceinline__ double2 operator()(int pt, int half) const
    {
        return __ldg(reinterpret_cast<const double2 *>(m + 4 * pt + 2 * half));
    }
};

// ------------------------------------------------------------------------------------------
// Element-local sweep (MODE_LAPLACE in place, MODE_SWEEP1 qtens -> tmp): warp-private TMA pipeline.
// grid = (elements, units); the 4 warps of a CTA share one element (one staged metric) and
// interleave its tiles.
// ------------------------------------------------------------------------------------------
constexpr int TILE_BYTES = TILE_SLABS * 128;

template <int STAGES, int WARPS>
constexpr int sweep_smem_bytes()
{
    // 1024 alignment slack + stages + mbarriers + metric + rspheremp
    return 1024 + WARPS * STAGES * TILE_BYTES + ((WARPS * STAGES * 8 + 15) / 16) * 16 + METRIC_DOUBLES * 8 + 16 * 8 + 64;
}

__device__ __forceinline__ void ldg256(const void *ptr, double &a, double &b, double &c, double &d)
{
    asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(ptr));
}
__device__ __forceinline__ double ldg64(const void *ptr)
{
    double a;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(a) : "l"(ptr));
    return a;
}

template <int MODE, int STAGES, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) sweep_kernel(const __grid_constant__ SweepParams p)
{
    constexpr int SB = TILE_BYTES;
    extern __shared__ unsigned char smem_raw[];
    // keep the pointer derived from smem_raw so that the compiler emits LDS/STS, not generic LD/ST
    unsigned char *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char *stages = smem;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + WARPS * STAGES * SB);
    double *metric_s = reinterpret_cast<double *>(smem + WARPS * STAGES * SB + ((WARPS * STAGES * 8 + 15) / 16) * 16);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int le = p.elem_list ? p.elem_list[blockIdx.x] : p.elem_first + (int)blockIdx.x;

    // ---- per-CTA staging of the element's folded metric
    for (int t = threadIdx.x; t < METRIC_DOUBLES; t += WARPS * 32) metric_s[t] = p.metric[(size_t)le * METRIC_DOUBLES + t];
    double *rsph_s = metric_s + METRIC_DOUBLES;
    constexpr bool S1 = MODE == MODE_SWEEP1 || MODE == MODE_SWEEP1_PRE;
    if (S1 && threadIdx.x < 16) rsph_s[threadIdx.x] = p.rsph[(size_t)le * 16 + threadIdx.x];
    // MODE_SWEEP1_PRE: the element's per-level factors [nlev][16], 16-byte chunk c of level k at chunk c ^ (k & 7) so that
    // lanes holding consecutive levels read their 128-byte rows without bank conflicts
    unsigned char *pre_s = reinterpret_cast<unsigned char *>(rsph_s + 16);
    if (MODE == MODE_SWEEP1_PRE) {
        const double2 *src = reinterpret_cast<const double2 *>(p.prescale + (size_t)le * p.nlev * 16);
        for (int t = threadIdx.x; t < p.nlev * 8; t += WARPS * 32) {
            const int k = t >> 3, c = t & 7;
            *reinterpret_cast<double2 *>(pre_s + k * 128 + ((c ^ (k & 7)) << 4)) = __ldg(src + t);
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(smem_u32(&bars[warp * STAGES + s]), 1);
        fence_barrier_init();
        fence_proxy_async();
    }
    __syncthreads();

    // ---- this warp's tiles: t0 + warp, t0 + warp + WARPS, ...
    const int ntiles = (p.kq + TILE_SLABS - 1) / TILE_SLABS;
    const int t0 = blockIdx.y * p.tiles_per_unit;
    const int t1 = min(t0 + p.tiles_per_unit, ntiles);
    const int my_n = (t1 - t0 > warp) ? (t1 - t0 - warp + WARPS - 1) / WARPS : 0;

    const uint32_t stage0 = smem_u32(stages + (size_t)warp * STAGES * SB);
    const uint32_t bar0 = smem_u32(&bars[warp * STAGES]);

    auto issue_load = [&](int n, int s) {
        const int kq0 = (t0 + warp + n * WARPS) * TILE_SLABS;
        const uint32_t dst = stage0 + s * SB, bar = bar0 + s * 8;
        mbar_expect_tx(bar, (uint32_t)TILE_BYTES);
        tma_load_3d(dst, &p.map_q, 0, kq0, le, bar);
    };

    if (lane == 0) {
        const int pre = my_n < STAGES ? my_n : STAGES;
        for (int n = 0; n < pre; ++n) issue_load(n, n);
    }

    const SmemMetric M{metric_s};
    const int sw = lane & 7; // SWIZZLE_128B: 16-byte chunk c of smem row r sits at chunk c ^ (r & 7)

    for (int n = 0; n < my_n; ++n) {
        const int s = n % STAGES;
        const int kq0 = (t0 + warp + n * WARPS) * TILE_SLABS;
        unsigned char *tile = stages + ((size_t)warp * STAGES + s) * SB;

        mbar_wait(bar0 + s * 8, (uint32_t)((n / STAGES) & 1));

        double v[16];
        unsigned char *row = tile + lane * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
            v[2 * c] = t.x;
            v[2 * c + 1] = t.y;
        }

        if (MODE == MODE_SWEEP1_PRE) { // Qtens = Qdp * (dpdiss/dp) of this slab's level (lanes past a ragged end compute on clipped zeros)
            const int k = (kq0 + lane) % p.nlev;
            const unsigned char *prow = pre_s + k * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const double2 r = *reinterpret_cast<const double2 *>(prow + ((c ^ (k & 7)) << 4));
                v[2 * c] *= r.x;
                v[2 * c + 1] *= r.y;
            }
        }
        laplace_slab(v, M);
        if (S1) {
            // DSS scaling folded into the producer: rspheremp is a property of the NODE (the same in every
            // element sharing it), so (sum_e q_e) * rspheremp == sum_e (q_e * rspheremp) up to rounding
            // and sweep 2 only has to add the neighbours' (already scaled) values
#pragma unroll
            for (int t = 0; t < 16; t += 2) {
                const double2 r = *reinterpret_cast<const double2 *>(rsph_s + t);
                v[t] *= r.x;
                v[t + 1] *= r.y;
            }
        }

#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<double2 *>(row + ((c ^ sw) << 4)) = make_double2(v[2 * c], v[2 * c + 1]);
        fence_proxy_async();
        __syncwarp();

        if (lane == 0) {
            tma_store_3d(S1 ? &p.map_tmp : &p.map_q, 0, kq0, le, stage0 + s * SB);
            bulk_commit();
            // refill the stage tile n-1 used, once its store has finished reading shared memory
            if (n >= 1 && n - 1 + STAGES < my_n) {
                bulk_wait_read<1>();
                issue_load(n - 1 + STAGES, (n - 1) % STAGES);
            }
        }
    }
    if (lane == 0) bulk_wait_read<0>(); // shared memory may be released once the stores have READ it; the writes complete on their own (kernel end flushes them)
}

// ------------------------------------------------------------------------------------------
// Sweep 2 of the full biharmonic (HOMME edgeVunpack + rspheremp + second laplace_sphere_wk):
//     tmp + the neighbours' coincident edge/corner values -> * rspheremp -> laplace -> qtens
//
// A CTA owns a W-E *strip* of up to WARPS consecutive elements (warp w <-> element first+w) and
// walks the same (k,q) tiles for all of them in lock step.  Each warp TMA-loads its own element's
// tile.  W/E edges between strip mates never leave the SM: a warp reads the facing column of its
// mate's tile from shared memory.  Every other contributor (S/N neighbours' rows, the strip ends'
// W/E columns, diagonal corner values, other ranks' halo rows) is read straight from the
// neighbour's slab in tmp / from the halo buffer with 256-bit or 64-bit global loads issued
// before the tile's mbarrier wait.  Those lines are the ones the neighbour's own CTA streams
// at (nearly) the same time -- strips are launched in element order, S/N neighbours are ne/4
// CTAs apart -- so they are served by the 126 MB L2 and DRAM sees each slab once.
// Corners are summed in ascending global point order (bit-identical for any rank count).
// ------------------------------------------------------------------------------------------
// what the strip kernel does after assembling the DSS sums
enum : int { SK_DSS = 0,   // nothing: edgeVunpack only (values were scaled by rspheremp by their producer)
             SK_LAP = 1,   // second laplace_sphere_wk
             SK_POST = 2 };// second Laplacian, then advance_hypervis_scalar's update fused in: the slab of Qdp is TMA-loaded into
                           // the tile's own stage while the Laplacian runs, out = limiter2d_zero(Qdp*spheremp - coef*lap)*rspheremp

// Qdp*spheremp - coef*Qtens with the CPU statement's roundings, limiter2d_zero in the CPU order (a thread owns the level)
__device__ __forceinline__ void hypervis_update_slab(double (&v)[16], const double (&a)[16], const double *mp, const double *rs, double coef, int limiter)
{
    double x[16];
#pragma unroll
    for (int t = 0; t < 16; ++t) x[t] = __dsub_rn(__dmul_rn(a[t], mp[t]), __dmul_rn(coef, v[t]));
    if (limiter) {
        double mass = 0.0;
#pragma unroll
        for (int t = 0; t < 16; ++t) mass = __dadd_rn(mass, x[t]);
        const bool neg = mass < 0.0;
        double added = 0.0;
#pragma unroll
        for (int t = 0; t < 16; ++t) {
            if (neg) x[t] = -x[t];
            if (x[t] < 0.0) x[t] = 0.0;
            else added = __dadd_rn(added, x[t]);
        }
        if (added > 0.0) {
            const double scale = fabs(mass) / added;
#pragma unroll
            for (int t = 0; t < 16; ++t) x[t] = __dmul_rn(x[t], scale);
        }
        if (neg) {
#pragma unroll
            for (int t = 0; t < 16; ++t) x[t] = -x[t];
        }
    }
#pragma unroll
    for (int t = 0; t < 16; ++t) v[t] = x[t] * rs[t];
}

struct WarpDss {                  // per warp in shared memory, built once per CTA (112 bytes)
    const unsigned char *edge_ptr[4]; // S,N,W,E: address of value 0 for slab 0
    const unsigned char *diag_ptr[4]; // corners (1,1),(4,1),(1,4),(4,4): address for slab 0, nullptr = none
    int32_t edge_stride[4];       // bytes between consecutive (k,q) slabs
    int32_t diag_stride[4];
    uint8_t edge_mode[4];         // EM_*
    uint8_t edge_rev[4];
    uint8_t corner_n[4];
    uint8_t corner_perm[4];
    // 1 = "face interior" element: S/N rows and W/E columns (global or strip mate) all present and
    // not reversed, four diagonal contributors, and the ascending-global-id summation order of a
    // row-major structured patch -> straight-line code without mode branches and selects
    int32_t fast;
    int32_t pad;
};
// corner_perm of a face-interior element with element ids row-major (e = je*ne + ie), 2 bits per
// summation slot (0 own, 1 S/N neighbour, 2 W/E neighbour, 3 diagonal):
//   SW: diag, S, W, own   SE: S, diag, own, E   NW: W, own, diag, N   NE: own, E, N, diag
constexpr int PERM_SW = 3 | (1 << 2) | (2 << 4) | (0 << 6), PERM_SE = 1 | (3 << 2) | (0 << 4) | (2 << 6);
constexpr int PERM_NW = 2 | (0 << 2) | (3 << 4) | (1 << 6), PERM_NE = 0 | (2 << 2) | (1 << 4) | (3 << 6);
enum : int { EM_NONE = 0, EM_ROW = 1, EM_COL = 2, EM_MATE_LEFT = 3, EM_MATE_RIGHT = 4 };

// Shared memory of the strip kernel, in order (offsets from a 1024-byte aligned base):
//   tiles   [warp][stage][4 KB]                        TMA destination, SWIZZLE_128B, results written in place
//   rowbuf  [set 2][warp][S,N][32 slabs][32 B]         facing rows of the S/N neighbours, TMA destination, SWIZZLE_32B
//   endbuf  [W end,E end][column: 4 x 512 B | diagonal 512 B | diagonal 512 B]   strip-end neighbours, TMA destination, single buffered
//   barriers: full [warp][stage] (tile + S/N rows + strip-end boxes of one tile), 2 spare slots per warp, cons [warp][2],
//             1 spare slot per warp, qdp_full [warp] (SK_POST)
//   metric [warp][64], endtma [6], spheremp + rspheremp [warp][32] (SK_POST)
// The generic path keeps its per-warp descriptors (WarpDss) in the rowbuf area.
constexpr int ROWBUF_BYTES = 2 * STRIP_MAX * 2 * 32 * 32;
constexpr int END_SIDE_BYTES = 2048 + 2 * 512;
constexpr int ENDBUF_BYTES = 2 * END_SIDE_BYTES;
constexpr int STRIP_EXTRA_BYTES = ROWBUF_BYTES + ENDBUF_BYTES;
static_assert(ROWBUF_BYTES >= STRIP_MAX * (int)sizeof(WarpDss) && STRIP_EXTRA_BYTES % 16 == 0, "the generic path's records share the staging area");
template <int STAGES, int WARPS>
constexpr int strip_smem_bytes()
{
    return 1024 + WARPS * STAGES * TILE_BYTES + STRIP_EXTRA_BYTES + WARPS * (STAGES + 6) * 8 + WARPS * METRIC_DOUBLES * 8 + 6 * 16 + WARPS * 32 * 8 + 64;
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, uint32_t src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async8(uint32_t dst, const void *src, uint32_t src_bytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
// the mbarrier gets one arrival from this thread once all its earlier cp.async have landed
__device__ __forceinline__ void cp_async_arrive(uint32_t bar)
{
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

// ---- regular path (StripReg): S/N neighbour rows stageable by TMA, free strip ends.  No CTA-wide barrier inside the
// tile loop: every warp runs its own pipeline.
//   * own tile: TMA, STAGES-deep ring, results written in place, TMA store (as in sweep 1);
//   * facing rows of the S/N neighbours: two TMA boxes {4 doubles, 32 slabs} per tile into rowbuf;
//   * strip ends: facing column (or row) and two diagonal corner values of the outside neighbours come by TMA too.
//     The neighbour's slab is viewed as {i,j,kq,elem}; a box {2,1,32 slabs} picks the 16-byte i-pair that holds one
//     value of 32 consecutive slabs (16 bytes is the TMA minimum): four of them for a facing column, one per
//     diagonal; a {4,32} row box when the neighbour faces with a row or sits in the halo.  (Gathering these with
//     per-lane cp.async cost 32 LSU wavefronts per instruction -- a third of the kernel's shared-memory pipe
//     traffic; a single swizzled {2,4,32} box is rejected by the hardware: swizzle modes fault when the inner box
//     is narrower than the swizzle span.)  endbuf is single buffered: its values are in registers a few
//     instructions into the tile, and the next tile's boxes are issued before the Laplacian, tracked by end_full;
//   * rowbuf is double buffered, refilled two tiles ahead and tracked by nbr_full (lane 0's expect_tx);
//   * mates: a warp reads the facing column of its mate's tile and the facing ends of the mate's S/N
//     neighbour rows (= its own diagonal contributors) straight from the mate's shared memory, after
//     waiting on the MATE's tile_full / nbr_full barriers, and then arrives on the mate's cons barrier
//     (parity of the tile).  The owner waits on its cons barrier before it overwrites the tile with its
//     results and before it refills the rowbuf set.  Arrivals depend only on loads having landed, never on
//     a mate's computation, so the protocol cannot dead-lock; and a mate can run at most one tile ahead
//     (it needs the owner's arrival for its own tile), so two cons barriers per warp are enough.
enum : int { ET_NONE = 0, ET_COL, ET_ROW, ET_HALOROW, ET_PT, ET_HALOPT };
struct EndTma {          // one staged source of a strip end (16 bytes)
    uint8_t kind;        // ET_*
    uint8_t hi;          // which double of the staged i-pair holds the value (ET_COL, ET_PT, ET_HALOPT)
    uint8_t rev, pad;    // column/row runs the other way
    int32_t c0, c1;      // first i (or first double of the row) / j of the box
    int32_t src;         // local element or halo slot
};
__device__ __forceinline__ uint32_t end_box_bytes(int kind)
{
    return kind == ET_COL ? 2048u : (kind == ET_ROW || kind == ET_HALOROW) ? 1024u : kind == ET_NONE ? 0u : 512u;
}

template <int STAGES, int WARPS, int KIND>
__device__ __forceinline__ void sweep2_regular(const SweepParams &p, unsigned char *stages, uint64_t *bars, double *metric_all, unsigned char *extra,
                                               int first, int cnt, const StripReg &rg)
{
    constexpr int SB = TILE_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool active = warp < cnt;
    const int le = first + warp;
    double *metric_s = metric_all + warp * METRIC_DOUBLES;
    unsigned char *rowbuf = extra;                        // [set][warp][2][32][32 B]
    unsigned char *endbuf = extra + ROWBUF_BYTES;         // [set][2][6][32][8 B]
    uint64_t *nbars = bars + WARPS * STAGES;              // (two spare slots per warp)
    uint64_t *cbars = nbars + WARPS * 2;                  // cons     [warp][2]
    uint64_t *ebars = cbars + WARPS * 2;                  // end_full [warp]
    // sources of the two strip ends, decoded once per CTA: [W column, SW, NW, E column, SE, NE]
    EndTma *endtma = reinterpret_cast<EndTma *>(metric_all + STRIP_MAX * METRIC_DOUBLES);

    const bool w_end = warp == 0, e_end = warp == cnt - 1;
    const int ntiles = (p.kq + TILE_SLABS - 1) / TILE_SLABS;
    const int t0 = blockIdx.y * p.tiles_per_unit;
    const int my_n = min(t0 + p.tiles_per_unit, ntiles) - t0;
    const uint32_t stage0 = smem_u32(stages + (size_t)warp * STAGES * SB);
    const uint32_t bar0 = smem_u32(&bars[warp * STAGES]);
    if (lane == 0) {
#pragma unroll
        // one barrier per tile in flight collects everything that tile needs: the tile itself, the S/N rows and, for an end
        // warp, the strip-end boxes -- three separate expect_tx arrivals (issued 3, 2 and 1 tiles ahead), one wait
        for (int s = 0; s < STAGES; ++s) mbar_init(smem_u32(&bars[warp * STAGES + s]), (w_end || e_end) ? 3u : 2u);
        mbar_init(smem_u32(&ebars[WARPS + warp]), 1); // qdp_full (SK_POST)
        const uint32_t nmates = (w_end ? 0u : 1u) + (e_end ? 0u : 1u);
        mbar_init(smem_u32(&cbars[warp * 2]), active && nmates ? 32u * nmates : 1u); // every lane of every mate arrives
        mbar_init(smem_u32(&cbars[warp * 2 + 1]), active && nmates ? 32u * nmates : 1u);
        fence_barrier_init();
        fence_proxy_async();
        // this warp's first tiles need nothing but its own barriers: get them on their way before the descriptor decode
        if (active) {
            const int pre = my_n < STAGES ? my_n : STAGES;
            for (int n = 0; n < pre; ++n) {
                mbar_expect_tx(bar0 + n * 8, (uint32_t)TILE_BYTES);
                tma_load_3d(stage0 + n * SB, &p.map_tmp, 0, (t0 + n) * TILE_SLABS, le, bar0 + n * 8);
            }
        }
    }
    double *hv_s = reinterpret_cast<double *>(endtma + 6) + warp * 32; // SK_POST: this element's spheremp, rspheremp
    const uint32_t qbar = smem_u32(&ebars[WARPS + warp]);
    if (KIND == SK_POST && active) hv_s[lane] = lane < 16 ? p.elem[(size_t)le * ELEM_DOUBLES + 64 + lane] : p.rsph[(size_t)le * 16 + lane - 16];
    if (active) { // the folded metric (512 B) arrives asynchronously, behind the first tile loads
        cp_async16(smem_u32(metric_s) + lane * 16, p.metric + (size_t)le * METRIC_DOUBLES + lane * 2, 16u);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (threadIdx.x < 6) {
        const int k = threadIdx.x, side = k / 3, which = k % 3; // which: 0 column, 1 S-side diagonal, 2 N-side diagonal
        EndTma e{ET_NONE, 0, 0, 0, 0, 0, 0};
        if (which == 0) {
            const int kind = rg.end_kind[side], row = rg.end_row[side];
            if (kind == SRC_TMPCOL) e = {ET_COL, (uint8_t)(row ? 1 : 0), 0, 0, row ? 2 : 0, 0, 0}; // column i=1: first of the pair (1,2); i=4: second of (3,4)
            else if (kind == SRC_TMP) e = {ET_ROW, 0, 0, 0, 4 * row, 0, 0};
            else e = {ET_HALOROW, 0, 0, 0, 0, 0, 0};
            e.rev = rg.end_rev[side];
            e.src = rg.end_src[side];
        } else {
            const int d = side * 2 + (which - 1);
            if (rg.diag_src[d] >= 0) {
                const int pt = rg.diag_pt[d];
                if (rg.diag_kind[d] == SRC_HALO) e = {ET_HALOPT, (uint8_t)(pt & 1), 0, 0, pt & 2, 0, 0};
                else e = {ET_PT, (uint8_t)(pt & 1), 0, 0, pt & 2, pt >> 2, 0};
                e.src = rg.diag_src[d];
            }
        }
        endtma[k] = e;
    }
    __syncthreads(); // the only CTA-wide barrier: mbarriers initialised, endtma written
    if (!active) return;

    const uint32_t cbar0 = smem_u32(&cbars[warp * 2]);
    const int swz = (lane >> 2) & 1; // SWIZZLE_32B: 16-byte chunk c of 32-byte row r sits at chunk c ^ ((r >> 2) & 1)
    const int elemS = rg.s_src[warp & 3], elemN = rg.n_src[warp & 3];
    const int rawS = rg.s_c0[warp & 3], rawN = rg.n_c0[warp & 3];
    const CUtensorMap *mapS = rawS == ROW_IN_HALO ? &p.map_halo : &p.map_rows, *mapN = rawN == ROW_IN_HALO ? &p.map_halo : &p.map_rows;
    const int c0S = rawS == ROW_IN_HALO ? 0 : rawS, c0N = rawN == ROW_IN_HALO ? 0 : rawN;
    const bool std_corners = rg.all_std != 0;
    uint32_t cn_pack = 0, cperm_pack = 0; // this member's four corner counts / summation orders, one byte each
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        cn_pack |= (uint32_t)rg.corner_n[warp & 3][c] << (8 * c);
        cperm_pack |= (uint32_t)rg.corner_perm[warp & 3][c] << (8 * c);
    }
    const uint64_t pol_out = l2_policy_evict_first(); // the output is never read again by this kernel

    auto issue_tile = [&](int n, int s) {
        const uint32_t dst = stage0 + s * SB, bar = bar0 + s * 8;
        mbar_expect_tx(bar, (uint32_t)TILE_BYTES);
        tma_load_3d(dst, &p.map_tmp, 0, (t0 + n) * TILE_SLABS, le, bar);
    };
    // lane 0: stage the S/N neighbours' facing rows of tile n into buffer set n & 1
    auto issue_nbr = [&](int n) {
        const int q = n & 1;
        const int kq0 = (t0 + n) * TILE_SLABS;
        const uint32_t nbar = bar0 + (n % STAGES) * 8;
        const uint32_t rb = smem_u32(rowbuf) + (uint32_t)((q * STRIP_MAX + warp) * 2048);
        mbar_expect_tx(nbar, 2048u);
        tma_load_3d(rb, mapS, c0S, kq0, elemS, nbar);        // the S neighbour's facing row (its N row in a face interior)
        tma_load_3d(rb + 1024, mapN, c0N, kq0, elemN, nbar); // the N neighbour's facing row
    };
    // lane 0 of an end warp: stage the outside neighbour's facing column and the two diagonal corner values of tile n
    const bool is_end = w_end || e_end;
    uint32_t end_bytes = 0;
    if (is_end) {
#pragma unroll
        for (int k = 0; k < 6; ++k)
            if (k < 3 ? w_end : e_end) end_bytes += end_box_bytes(endtma[k].kind);
    }
    auto issue_end = [&](int n) {
        const int kq0 = (t0 + n) * TILE_SLABS;
        const uint32_t ebar = bar0 + (n % STAGES) * 8;
        mbar_expect_tx(ebar, end_bytes);
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            if (side == 0 ? !w_end : !e_end) continue;
#pragma unroll
            for (int which = 0; which < 3; ++which) {
                const EndTma e = endtma[3 * side + which];
                const uint32_t dst = smem_u32(endbuf) + (uint32_t)(side * END_SIDE_BYTES + (which ? 1536 + 512 * which : 0));
                switch (e.kind) {
                case ET_COL:
#pragma unroll
                    for (int j = 0; j < 4; ++j) tma_load_4d(dst + j * 512, &p.map_pt, e.c0, j, kq0, e.src, ebar);
                    break;
                case ET_ROW: tma_load_3d(dst, &p.map_rows, e.c0, kq0, e.src, ebar); break;
                case ET_HALOROW: tma_load_3d(dst, &p.map_halo, 0, kq0, e.src, ebar); break;
                case ET_PT: tma_load_4d(dst, &p.map_pt, e.c0, e.c1, kq0, e.src, ebar); break;
                case ET_HALOPT: tma_load_3d(dst, &p.map_halo_pt, e.c0, kq0, e.src, ebar); break;
                default: break; // no diagonal contributor (cube corner)
                }
            }
        }
    };
    // all lanes of an end warp: the four values along the facing edge (own orientation) and the two diagonals
    auto read_end = [&](int side, double (&col)[4], double &d0, double &d1) {
        const unsigned char *eb = endbuf + side * END_SIDE_BYTES;
        const EndTma c = endtma[3 * side];
        const int rm = c.rev ? 3 : 0; // position t of this element's edge = position t ^ rm of the neighbour's
        if (c.kind == ET_COL) { // [j][slab][2 doubles]
            const unsigned char *b = eb + lane * 16 + c.hi * 8;
#pragma unroll
            for (int t = 0; t < 4; ++t) col[t] = *reinterpret_cast<const double *>(b + ((t ^ rm) << 9));
        } else { // [slab][4 doubles], SWIZZLE_32B
            const unsigned char *b = eb + lane * 32;
            const int x = ((lane >> 2) & 1) ^ (rm >> 1), y = rm & 1;
#pragma unroll
            for (int t = 0; t < 4; ++t) col[t] = *reinterpret_cast<const double *>(b + (((t >> 1) ^ x) << 4) + (((t & 1) ^ y) << 3));
        }
        // [slab][2 doubles]; an absent diagonal (cube corner) leaves the box unwritten and the value unused
        d0 = *reinterpret_cast<const double *>(eb + 2048 + lane * 16 + endtma[3 * side + 1].hi * 8);
        d1 = *reinterpret_cast<const double *>(eb + 2560 + lane * 16 + endtma[3 * side + 2].hi * 8);
    };

    if (lane == 0) { // (the first tile loads were issued before the decode)
        if (my_n > 0) issue_nbr(0);
        if (my_n > 1) issue_nbr(1);
        if (is_end && my_n > 0) issue_end(0);
    }

    const SmemMetric M{metric_s};
    const int sw = lane & 7;

    for (int n = 0; n < my_n; ++n) {
        const int s = n % STAGES, q = n & 1;
        const int kq0 = (t0 + n) * TILE_SLABS;
        const uint32_t parity = (uint32_t)((n / STAGES) & 1), qparity = (uint32_t)((n >> 1) & 1);
        unsigned char *row = stages + ((size_t)warp * STAGES + s) * SB + lane * 128;
        const unsigned char *rb = rowbuf + (size_t)((q * STRIP_MAX + warp) * 2048) + lane * 32;
        double v[16], nb[4][4], dg[4];

        mbar_wait(bar0 + s * 8, parity);
        if (n == 0) { // the metric staged by this warp's own cp.async group
            asm volatile("cp.async.wait_all;" ::: "memory");
            __syncwarp();
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
            v[2 * c] = t.x;
            v[2 * c + 1] = t.y;
        }
        {
            const double2 a = *reinterpret_cast<const double2 *>(rb + (swz << 4)), b = *reinterpret_cast<const double2 *>(rb + ((1 ^ swz) << 4));
            nb[EDGE_S][0] = a.x; nb[EDGE_S][1] = a.y; nb[EDGE_S][2] = b.x; nb[EDGE_S][3] = b.y;
            const double2 c = *reinterpret_cast<const double2 *>(rb + 1024 + (swz << 4)), d = *reinterpret_cast<const double2 *>(rb + 1024 + ((1 ^ swz) << 4));
            nb[EDGE_N][0] = c.x; nb[EDGE_N][1] = c.y; nb[EDGE_N][2] = d.x; nb[EDGE_N][3] = d.y;
        }
        // ---- W side: outside neighbour (strip end) or the left mate's shared memory
        if (w_end) {
            read_end(0, nb[EDGE_W], dg[0], dg[2]);
        } else { // mate's E column = points 3,7,11,15 = upper halves of chunks 1,3,5,7; east ends of its S/N neighbour rows
            mbar_wait(bar0 - STAGES * 8 + s * 8, parity);
            const unsigned char *mrow = row - STAGES * SB;
#pragma unroll
            for (int t = 0; t < 4; ++t) nb[EDGE_W][t] = *reinterpret_cast<const double *>(mrow + (((2 * t + 1) ^ sw) << 4) + 8);
            dg[0] = *reinterpret_cast<const double *>(rb - 2048 + ((1 ^ swz) << 4) + 8);        // mate's S-row value 3
            dg[2] = *reinterpret_cast<const double *>(rb - 2048 + 1024 + ((1 ^ swz) << 4) + 8); // mate's N-row value 3
        }
        // ---- E side
        if (e_end) {
            read_end(1, nb[EDGE_E], dg[1], dg[3]);
        } else { // mate's W column = points 0,4,8,12 = lower halves of chunks 0,2,4,6; west ends of its S/N neighbour rows
            mbar_wait(bar0 + STAGES * 8 + s * 8, parity);
            const unsigned char *mrow = row + STAGES * SB;
#pragma unroll
            for (int t = 0; t < 4; ++t) nb[EDGE_E][t] = *reinterpret_cast<const double *>(mrow + (((2 * t) ^ sw) << 4));
            dg[1] = *reinterpret_cast<const double *>(rb + 2048 + (swz << 4));        // mate's S-row value 0
            dg[3] = *reinterpret_cast<const double *>(rb + 2048 + 1024 + (swz << 4)); // mate's N-row value 0
        }

        // ---- DSS (edgeVunpack): mid-edge points, then corners in ascending global point order.
        //      (values were scaled by rspheremp in sweep 1)
        v[1] += nb[EDGE_S][1]; v[2] += nb[EDGE_S][2];
        v[13] += nb[EDGE_N][1]; v[14] += nb[EDGE_N][2];
        v[4] += nb[EDGE_W][1]; v[8] += nb[EDGE_W][2];
        v[7] += nb[EDGE_E][1]; v[11] += nb[EDGE_E][2];
        if (std_corners) {
            v[0] = ((dg[0] + nb[EDGE_S][0]) + nb[EDGE_W][0]) + v[0];       // PERM_SW
            v[3] = ((nb[EDGE_S][3] + dg[1]) + v[3]) + nb[EDGE_E][0];       // PERM_SE
            v[12] = ((nb[EDGE_W][3] + v[12]) + dg[2]) + nb[EDGE_N][0];     // PERM_NW
            v[15] = ((v[15] + nb[EDGE_E][3]) + nb[EDGE_N][3]) + dg[3];     // PERM_NE
        } else { // face edges, cube corners, partition cuts: up to four contributors in the plan's order
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int pt = c == 0 ? 0 : (c == 1 ? 3 : (c == 2 ? 12 : 15));
                const double cv0 = v[pt];
                const double cv1 = nb[c < 2 ? EDGE_S : EDGE_N][(c & 1) ? 3 : 0];
                const double cv2 = nb[(c & 1) ? EDGE_E : EDGE_W][c < 2 ? 0 : 3];
                const double cv3 = dg[c];
                const int perm = (cperm_pack >> (8 * c)) & 255, cn = (cn_pack >> (8 * c)) & 255;
                auto pick = [&](int k) {
                    const int w = (perm >> (2 * k)) & 3;
                    return w == 0 ? cv0 : (w == 1 ? cv1 : (w == 2 ? cv2 : cv3));
                };
                double sum = pick(0);
                if (cn > 1) sum = sum + pick(1);
                if (cn > 2) sum = sum + pick(2);
                if (cn > 3) sum = sum + pick(3);
                v[pt] = sum;
            }
        }

        // the mates' values have been consumed (the sums above depend on them): tell the mates.
        // The asm operand ties the arrival to the corner sums so that it cannot be scheduled above the loads.
        if (!w_end) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(cbar0 - 16 + q * 8), "d"(v[0]), "d"(v[12]), "d"(v[4]), "d"(v[8]) : "memory");
        if (!e_end) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(cbar0 + 16 + q * 8), "d"(v[3]), "d"(v[15]), "d"(v[7]), "d"(v[11]) : "memory");
        if (is_end) {
            // endbuf has been consumed by every lane (the sums above depend on its values): refill it for the next tile now,
            // so that the boxes travel while this tile's Laplacian runs
            asm volatile("" ::"d"(v[0]), "d"(v[12]), "d"(v[4]), "d"(v[8]), "d"(v[3]), "d"(v[15]), "d"(v[7]), "d"(v[11]) : "memory");
            __syncwarp();
            if (lane == 0 && n + 1 < my_n) issue_end(n + 1);
        }

        if (KIND == SK_POST) {
            // This warp holds the tile in registers and, once cons fires, the mates have read its columns: the stage is free
            // for the slab of Qdp, which travels while the Laplacian runs.
            if (cnt > 1) mbar_wait(cbar0 + q * 8, qparity);
            asm volatile("" ::"d"(v[5]), "d"(v[6]), "d"(v[9]), "d"(v[10]) : "memory");
            __syncwarp();
            if (lane == 0) {
                mbar_expect_tx(qbar, (uint32_t)TILE_BYTES);
                tma_load_3d(stage0 + s * SB, &p.map_q, 0, kq0, le, qbar);
            }
        }
        if (KIND != SK_DSS) laplace_slab(v, M); // MODE_DSS stops at the assembled (and already rspheremp-scaled) values

        if (KIND == SK_POST) {
            mbar_wait(qbar, (uint32_t)(n & 1));
            double a[16];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
                a[2 * c] = t.x;
                a[2 * c + 1] = t.y;
            }
            hypervis_update_slab(v, a, hv_s, hv_s + 16, p.hv_coef, p.hv_limiter);
        } else if (cnt > 1) {
            // both mates must have read this tile's columns and this rowbuf set before they are overwritten
            mbar_wait(cbar0 + q * 8, qparity);
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<double2 *>(row + ((c ^ sw) << 4)) = make_double2(v[2 * c], v[2 * c + 1]);
        fence_proxy_async();
        __syncwarp(); // all lanes' results are in the tile; all lanes' reads of rowbuf/endbuf set q have been consumed

        if (lane == 0) {
            tma_store_3d_hint(&p.map_out, 0, kq0, le, stage0 + s * SB, pol_out);
            bulk_commit();
            if (n >= 1 && n - 1 + STAGES < my_n) {
                bulk_wait_read<1>();
                issue_tile(n - 1 + STAGES, (n - 1) % STAGES);
            }
            // rowbuf set q is free again (own reads and the mates', see the cons wait above): stage tile n+2's rows into it
            if (n + 2 < my_n) issue_nbr(n + 2);
        }
    }
    if (lane == 0) bulk_wait_read<0>(); // shared memory may be released once the stores have READ it; the writes complete on their own (kernel end flushes them)
}

// ---- generic path: any connectivity the plan accepts (face edges, cube corners, halo rows, reversed edges)
template <int STAGES, int WARPS, int KIND>
__device__ __forceinline__ void sweep2_generic(const SweepParams &p, unsigned char *stages, uint64_t *bars, double *metric_all, unsigned char *extra,
                                               int first, int cnt)
{
    constexpr int SB = TILE_BYTES;
    WarpDss *wd_all = reinterpret_cast<WarpDss *>(extra);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool active = warp < cnt;
    const int le = first + warp;
    double *metric_s = metric_all + warp * METRIC_DOUBLES;
    WarpDss &wd = wd_all[warp];

    if (active) {
        metric_s[lane] = p.metric[(size_t)le * METRIC_DOUBLES + lane];
        metric_s[lane + 32] = p.metric[(size_t)le * METRIC_DOUBLES + lane + 32];
        if (lane < 4) {
            // lane e decodes edge e and corner e of the element's DSS descriptor
            const ElemDss &d = p.dss[le];
            const int e = lane;
            const size_t kq = (size_t)p.kq;
            const int src = d.edge_src[e];
            const int kind = d.edge_kind[e], row = d.edge_row[e], rev = d.edge_rev[e];
            const unsigned char *ptr = nullptr;
            int stride = 0, mode = EM_NONE;
            if (src >= 0) {
                if (kind == SRC_TMP) { ptr = reinterpret_cast<const unsigned char *>(p.tmp + ((size_t)src * kq * 16 + row * 4)); stride = 128; mode = EM_ROW; }
                else if (kind == SRC_TMPCOL) {
                    if (e == EDGE_W && !rev && row == 1 && warp > 0 && src == le - 1) mode = EM_MATE_LEFT;
                    else if (e == EDGE_E && !rev && row == 0 && warp + 1 < cnt && src == le + 1) mode = EM_MATE_RIGHT;
                    else { ptr = reinterpret_cast<const unsigned char *>(p.tmp + ((size_t)src * kq * 16 + (row ? 3 : 0))); stride = 128; mode = EM_COL; }
                } else { ptr = reinterpret_cast<const unsigned char *>(p.halo + (size_t)src * kq * 4); stride = 32; mode = EM_ROW; }
            }
            wd.edge_ptr[e] = ptr;
            wd.edge_stride[e] = stride;
            wd.edge_mode[e] = (uint8_t)mode;
            wd.edge_rev[e] = (uint8_t)rev;
            const int dsrc = d.diag_src[e];
            const unsigned char *dptr = nullptr;
            int dstride = 0;
            if (dsrc >= 0) {
                if (d.diag_kind[e] == SRC_TMPCOL) { dptr = reinterpret_cast<const unsigned char *>(p.tmp + ((size_t)dsrc * kq * 16 + (d.diag_row[e] ? 3 : 0) + 4 * d.diag_pos[e])); dstride = 128; }
                else { dptr = reinterpret_cast<const unsigned char *>(p.halo + (size_t)dsrc * kq * 4 + d.diag_pos[e]); dstride = 32; }
            }
            wd.diag_ptr[e] = dptr;
            wd.diag_stride[e] = dstride;
            wd.corner_n[e] = d.corner_n[e];
            wd.corner_perm[e] = d.corner_perm[e];
            const int std_perm = e == 0 ? PERM_SW : (e == 1 ? PERM_SE : (e == 2 ? PERM_NW : PERM_NE));
            const bool edge_ok = !rev && (e < 2 ? mode == EM_ROW : (mode == EM_COL || mode == EM_MATE_LEFT || mode == EM_MATE_RIGHT));
            const bool ok = edge_ok && dptr != nullptr && d.corner_n[e] == 4 && d.corner_perm[e] == std_perm;
            const unsigned all_ok = __ballot_sync(0xfu, ok);
            if (lane == 0) wd.fast = all_ok == 0xfu ? 1 : 0;
        }
    }
    // Strip members are linked E->W by construction (bwk_plan.cpp), so warp w reads the facing column of
    // the tiles of warps w-1 and w+1 and they read its own: same consumed-barrier protocol as the regular path.
    const bool has_left = active && warp > 0, has_right = active && warp < cnt - 1;
    uint64_t *cbars = bars + WARPS * STAGES + WARPS * 2; // cons [warp][2]
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(smem_u32(&bars[warp * STAGES + s]), 1);
        const uint32_t nmates = (has_left ? 1u : 0u) + (has_right ? 1u : 0u);
        mbar_init(smem_u32(&cbars[warp * 2]), nmates ? 32u * nmates : 1u);
        mbar_init(smem_u32(&cbars[warp * 2 + 1]), nmates ? 32u * nmates : 1u);
        mbar_init(smem_u32(&cbars[WARPS * 2 + WARPS + warp]), 1); // qdp_full (SK_POST), same slot as in the regular path
        fence_barrier_init();
        fence_proxy_async();
    }
    // SK_POST: this element's spheremp, rspheremp (behind the metric and the six strip-end records of the regular path)
    double *hv_s = reinterpret_cast<double *>(reinterpret_cast<unsigned char *>(metric_all + STRIP_MAX * METRIC_DOUBLES) + 6 * 16) + warp * 32;
    const uint32_t qbar = smem_u32(&cbars[WARPS * 2 + WARPS + warp]);
    if (KIND == SK_POST && active) hv_s[lane] = lane < 16 ? p.elem[(size_t)le * ELEM_DOUBLES + 64 + lane] : p.rsph[(size_t)le * 16 + lane - 16];
    __syncthreads(); // the only CTA-wide barrier
    if (!active) return;
    if ((has_left && wd.edge_mode[EDGE_W] != EM_MATE_LEFT) || (has_right && wd.edge_mode[EDGE_E] != EM_MATE_RIGHT)) __trap(); // plan/kernel disagree
    const uint32_t cbar0 = smem_u32(&cbars[warp * 2]);

    // ---- all warps walk the same tiles t0 .. t1-1, each for its own element
    const int ntiles = (p.kq + TILE_SLABS - 1) / TILE_SLABS;
    const int t0 = blockIdx.y * p.tiles_per_unit;
    const int my_n = min(t0 + p.tiles_per_unit, ntiles) - t0;

    const uint32_t stage0 = smem_u32(stages + (size_t)warp * STAGES * SB);
    const uint32_t bar0 = smem_u32(&bars[warp * STAGES]);

    auto issue_load = [&](int n, int s) {
        const uint32_t dst = stage0 + s * SB, bar = bar0 + s * 8;
        mbar_expect_tx(bar, (uint32_t)TILE_BYTES);
        tma_load_3d(dst, &p.map_tmp, 0, (t0 + n) * TILE_SLABS, le, bar);
    };
    if (lane == 0) {
        const int pre = my_n < STAGES ? my_n : STAGES;
        for (int n = 0; n < pre; ++n) issue_load(n, n);
    }

    const SmemMetric M{metric_s};
    const int sw = lane & 7;

    for (int n = 0; n < my_n; ++n) {
        const int s = n % STAGES;
        const int kq0 = (t0 + n) * TILE_SLABS;
        const uint32_t parity = (uint32_t)((n / STAGES) & 1);
        unsigned char *row = stages + ((size_t)warp * STAGES + s) * SB + lane * 128;
        double v[16], nb[4][4], dg[4];

        const int q = n & 1;
        const uint32_t qparity = (uint32_t)((n >> 1) & 1);
        const bool fast = wd.fast != 0;
        {
            // ---- contributors that live in global memory: issue the loads before the tile wait
            const bool live = kq0 + lane < p.kq;
            const size_t slab = (size_t)(kq0 + lane);
            if (fast) {
                // out-of-range lanes of a ragged last tile re-read slab 0 (valid memory, results clipped by the TMA store)
                const size_t sl = live ? slab : 0;
                ldg256(wd.edge_ptr[EDGE_S] + sl * (size_t)wd.edge_stride[EDGE_S], nb[EDGE_S][0], nb[EDGE_S][1], nb[EDGE_S][2], nb[EDGE_S][3]);
                ldg256(wd.edge_ptr[EDGE_N] + sl * (size_t)wd.edge_stride[EDGE_N], nb[EDGE_N][0], nb[EDGE_N][1], nb[EDGE_N][2], nb[EDGE_N][3]);
#pragma unroll
                for (int c = 0; c < 4; ++c) dg[c] = ldg64(wd.diag_ptr[c] + sl * (size_t)wd.diag_stride[c]);
                if (wd.edge_mode[EDGE_W] == EM_COL) {
                    const unsigned char *q = wd.edge_ptr[EDGE_W] + sl * 128;
#pragma unroll
                    for (int t = 0; t < 4; ++t) nb[EDGE_W][t] = ldg64(q + 32 * t);
                }
                if (wd.edge_mode[EDGE_E] == EM_COL) {
                    const unsigned char *q = wd.edge_ptr[EDGE_E] + sl * 128;
#pragma unroll
                    for (int t = 0; t < 4; ++t) nb[EDGE_E][t] = ldg64(q + 32 * t);
                }
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    nb[e][0] = nb[e][1] = nb[e][2] = nb[e][3] = 0.0;
                    const int mode = wd.edge_mode[e];
                    if (live && mode == EM_ROW) ldg256(wd.edge_ptr[e] + slab * (size_t)wd.edge_stride[e], nb[e][0], nb[e][1], nb[e][2], nb[e][3]);
                    else if (live && mode == EM_COL) {
                        const unsigned char *q = wd.edge_ptr[e] + slab * 128;
                        nb[e][0] = ldg64(q);
                        nb[e][1] = ldg64(q + 32);
                        nb[e][2] = ldg64(q + 64);
                        nb[e][3] = ldg64(q + 96);
                    }
                    dg[e] = 0.0;
                    const unsigned char *dp = wd.diag_ptr[e];
                    if (live && dp != nullptr) dg[e] = ldg64(dp + slab * (size_t)wd.diag_stride[e]);
                }
            }

            mbar_wait(bar0 + s * 8, parity);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
                v[2 * c] = t.x;
                v[2 * c + 1] = t.y;
            }
            // ---- strip mates: the facing column of the mate's tile, straight from shared memory
            if (wd.edge_mode[EDGE_W] == EM_MATE_LEFT) { // mate's E column: points 3,7,11,15 = upper half of chunks 1,3,5,7
                mbar_wait(bar0 - STAGES * 8 + s * 8, parity);
                const unsigned char *mrow = row - STAGES * SB;
#pragma unroll
                for (int t = 0; t < 4; ++t) nb[EDGE_W][t] = *reinterpret_cast<const double *>(mrow + (((2 * t + 1) ^ sw) << 4) + 8);
            }
            if (wd.edge_mode[EDGE_E] == EM_MATE_RIGHT) { // mate's W column: points 0,4,8,12 = lower half of chunks 0,2,4,6
                mbar_wait(bar0 + STAGES * 8 + s * 8, parity);
                const unsigned char *mrow = row + STAGES * SB;
#pragma unroll
                for (int t = 0; t < 4; ++t) nb[EDGE_E][t] = *reinterpret_cast<const double *>(mrow + (((2 * t) ^ sw) << 4));
            }
        }

        // ---- DSS: add the neighbours' coincident values (edgeVunpack; already scaled by rspheremp in sweep 1)
        if (fast) {
            v[1] += nb[EDGE_S][1]; v[2] += nb[EDGE_S][2];
            v[13] += nb[EDGE_N][1]; v[14] += nb[EDGE_N][2];
            v[4] += nb[EDGE_W][1]; v[8] += nb[EDGE_W][2];
            v[7] += nb[EDGE_E][1]; v[11] += nb[EDGE_E][2];
            // corners in ascending global point order of a face-interior element (PERM_SW/SE/NW/NE)
            v[0] = ((dg[0] + nb[EDGE_S][0]) + nb[EDGE_W][0]) + v[0];
            v[3] = ((nb[EDGE_S][3] + dg[1]) + v[3]) + nb[EDGE_E][0];
            v[12] = ((nb[EDGE_W][3] + v[12]) + dg[2]) + nb[EDGE_N][0];
            v[15] = ((v[15] + nb[EDGE_E][3]) + nb[EDGE_N][3]) + dg[3];
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (wd.edge_rev[e]) { // neighbour's edge runs the other way
                    double t = nb[e][0]; nb[e][0] = nb[e][3]; nb[e][3] = t;
                    t = nb[e][1]; nb[e][1] = nb[e][2]; nb[e][2] = t;
                }
            }
            // mid-edge points: exactly two contributors
            if (wd.edge_mode[EDGE_S] != EM_NONE) { v[1] += nb[EDGE_S][1]; v[2] += nb[EDGE_S][2]; }
            if (wd.edge_mode[EDGE_N] != EM_NONE) { v[13] += nb[EDGE_N][1]; v[14] += nb[EDGE_N][2]; }
            if (wd.edge_mode[EDGE_W] != EM_NONE) { v[4] += nb[EDGE_W][1]; v[8] += nb[EDGE_W][2]; }
            if (wd.edge_mode[EDGE_E] != EM_NONE) { v[7] += nb[EDGE_E][1]; v[11] += nb[EDGE_E][2]; }
            // corners: up to four contributors, summed in ascending global point order
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int pt = c == 0 ? 0 : (c == 1 ? 3 : (c == 2 ? 12 : 15));
                const double cv0 = v[pt];
                const double cv1 = nb[c < 2 ? EDGE_S : EDGE_N][(c & 1) ? 3 : 0];
                const double cv2 = nb[(c & 1) ? EDGE_E : EDGE_W][c < 2 ? 0 : 3];
                const double cv3 = dg[c];
                const int perm = wd.corner_perm[c], cn = wd.corner_n[c];
                auto pick = [&](int k) {
                    const int w = (perm >> (2 * k)) & 3;
                    return w == 0 ? cv0 : (w == 1 ? cv1 : (w == 2 ? cv2 : cv3));
                };
                double sum = pick(0);
                if (cn > 1) sum = sum + pick(1);
                if (cn > 2) sum = sum + pick(2);
                if (cn > 3) sum = sum + pick(3);
                v[pt] = sum;
            }
        }
        // the mates' columns have been consumed (the sums above depend on them): tell the mates
        if (has_left) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(cbar0 - 16 + q * 8), "d"(v[0]), "d"(v[12]), "d"(v[4]), "d"(v[8]) : "memory");
        if (has_right) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(cbar0 + 16 + q * 8), "d"(v[3]), "d"(v[15]), "d"(v[7]), "d"(v[11]) : "memory");

        if (KIND == SK_POST) { // as in the regular path: the stage takes the slab of Qdp while the Laplacian runs
            if (has_left || has_right) mbar_wait(cbar0 + q * 8, qparity);
            asm volatile("" ::"d"(v[5]), "d"(v[6]), "d"(v[9]), "d"(v[10]) : "memory");
            __syncwarp();
            if (lane == 0) {
                mbar_expect_tx(qbar, (uint32_t)TILE_BYTES);
                tma_load_3d(stage0 + s * SB, &p.map_q, 0, kq0, le, qbar);
            }
        }
        if (KIND != SK_DSS) laplace_slab(v, M); // (the rspheremp scaling was applied by sweep 1)

        if (KIND == SK_POST) {
            mbar_wait(qbar, (uint32_t)(n & 1));
            double a[16];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
                a[2 * c] = t.x;
                a[2 * c + 1] = t.y;
            }
            hypervis_update_slab(v, a, hv_s, hv_s + 16, p.hv_coef, p.hv_limiter);
        } else if (has_left || has_right) {
            // both mates must have read this tile's columns before it is overwritten in place
            mbar_wait(cbar0 + q * 8, qparity);
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<double2 *>(row + ((c ^ sw) << 4)) = make_double2(v[2 * c], v[2 * c + 1]);
        fence_proxy_async();
        __syncwarp();

        if (lane == 0) {
            tma_store_3d_hint(&p.map_out, 0, kq0, le, stage0 + s * SB, l2_policy_evict_first());
            bulk_commit();
            // refill the stage tile n-1 used: its store has read shared memory, and the mates
            // had consumed it before this warp overwrote it (cons wait of iteration n-1)
            if (n >= 1 && n - 1 + STAGES < my_n) {
                bulk_wait_read<1>();
                issue_load(n - 1 + STAGES, (n - 1) % STAGES);
            }
        }
    }
    if (lane == 0) bulk_wait_read<0>(); // shared memory may be released once the stores have READ it; the writes complete on their own (kernel end flushes them)
}

template <int STAGES, int WARPS, int MINB, int KIND>
__global__ void __launch_bounds__(WARPS * 32, MINB) sweep2_strip_kernel(const __grid_constant__ SweepParams p)
{
    static_assert(WARPS == STRIP_MAX, "one warp per strip element");
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char *stages = smem;                                  // tile ring, 1024-byte aligned (SWIZZLE_128B)
    unsigned char *extra = smem + WARPS * STAGES * TILE_BYTES;     // staging buffers (1024-aligned, SWIZZLE_32B rows) or WarpDss records
    uint64_t *bars = reinterpret_cast<uint64_t *>(extra + STRIP_EXTRA_BYTES);
    double *metric_all = reinterpret_cast<double *>(bars + WARPS * (STAGES + 6)); // after tile_full, nbr_full, cons, end_full, qdp_full

    const int32_t strip = p.strips[blockIdx.x];
    const int first = strip >> 3, cnt = (strip & 7) + 1;
    const StripReg &rg = p.strip_reg[blockIdx.x];
    if (rg.regular) sweep2_regular<STAGES, WARPS, KIND>(p, stages, bars, metric_all, extra, first, cnt, rg);
    else sweep2_generic<STAGES, WARPS, KIND>(p, stages, bars, metric_all, extra, first, cnt);
}

// ------------------------------------------------------------------------------------------
// Persistent form of the regular strip sweep (round 2).  Same per-tile protocol as sweep2_regular, but
//   * the grid is one wave (3 CTAs per SM) and a CTA walks work items i = blockIdx.x, + gridDim.x, ...;
//     item i = (unit i / nstrips, strip i % nstrips), a unit is only 3-5 tiles of the (k,q) axis.  All CTAs do the same
//     work per item, so the CTAs holding a strip and its S/N neighbour strips (30 items apart at ne120) stay within a few
//     tiles of each other for the whole launch: the neighbours' rows are read while the lines are still in L2
//     (short units gave 1.11x DRAM reads against 1.23-1.29x at 15 tiles in round 1, but paid 2.7 us of pipeline
//     fill/drain per CTA -- here the pipeline never drains);
//   * the TMA ring, the S/N row buffers, the strip-end boxes and every mbarrier run CONTINUOUSLY across items: lane 0
//     looks 1 (strip ends) and 2 (tile, rows) tiles ahead, into the next item when the current one runs out.  Items are
//     described by a small per-warp record (PDesc, EndTma) decoded from the strip's StripReg into one of two slots one
//     tile before the look-ahead first needs it; the folded metric is double buffered the same way (cp.async);
//   * only full regular strips (4 members) are taken, so every barrier's arrival count is a constant of the warp; the
//     host sends the other strips (face-edge strips with reversed rows, short strips at row ends) to sweep2_strip_kernel.
// Shared memory is carved to the byte: 76,640 B per CTA, three CTAs per SM.
// ------------------------------------------------------------------------------------------
struct PDesc {             // one work item as one warp sees it (64 bytes)
    int32_t le;            // the warp's element
    int32_t tile0, ntile;  // first tile of the item's unit, tiles in it (>= 3)
    int32_t elemS, elemN;  // sources of the S / N neighbour rows (local element or halo slot)
    uint32_t sel;          // c0S | c0N << 8 | S row in halo << 16 | N row in halo << 17 | all_std << 18
    uint32_t cn_pack, cperm_pack;
    uint32_t end_bytes;    // bytes of the strip-end boxes of one tile (end warps)
    int32_t pad[7];
};
static_assert(sizeof(PDesc) == 64 && sizeof(EndTma) == 16, "descriptor sizes are part of the shared-memory budget");
static_assert(offsetof(StripReg, s_src) == 4 && offsetof(StripReg, n_src) == 20 && offsetof(StripReg, s_c0) == 36 && offsetof(StripReg, n_c0) == 40 &&
              offsetof(StripReg, corner_n) == 44 && offsetof(StripReg, corner_perm) == 60 && offsetof(StripReg, end_src) == 76 &&
              offsetof(StripReg, end_kind) == 84 && offsetof(StripReg, end_row) == 86 && offsetof(StripReg, end_rev) == 88 &&
              offsetof(StripReg, all_std) == 90 && offsetof(StripReg, diag_src) == 92 && offsetof(StripReg, diag_kind) == 108 &&
              offsetof(StripReg, diag_pt) == 112 && offsetof(StripReg, first) == 116, "the persistent sweep decodes StripReg by 32-bit word");

template <int STAGES, int WARPS>
constexpr int persist_smem_bytes()
{
    return WARPS * STAGES * TILE_BYTES + STRIP_EXTRA_BYTES + 2 * WARPS * METRIC_DOUBLES * 8 + 2 * WARPS * (int)sizeof(PDesc) + 12 * (int)sizeof(EndTma) +
           WARPS * (STAGES + 2) * 8 + 2 * 8 + 16; // + the work queue's two mbarriers and four ints
}

template <int STAGES, int WARPS, int KIND>
__global__ void __launch_bounds__(WARPS * 32, 3) sweep2_persist_kernel(const __grid_constant__ SweepParams p)
{
    static_assert(WARPS == STRIP_MAX && KIND != SK_POST, "full strips, DSS or DSS + Laplacian");
    constexpr int SB = TILE_BYTES;
    extern __shared__ __align__(1024) unsigned char psm[];
    unsigned char *stages = psm;                                   // [warp][stage][4 KB], SWIZZLE_128B (needs the 1024-byte base)
    unsigned char *rowbuf = stages + WARPS * STAGES * SB;          // [set 2][warp][S,N][32][32 B]
    unsigned char *endbuf = rowbuf + ROWBUF_BYTES;                 // [W,E][column 2048 | diagonal 512 | diagonal 512]
    double *metric_all = reinterpret_cast<double *>(endbuf + ENDBUF_BYTES);           // [slot 2][warp][64]
    PDesc *desc_all = reinterpret_cast<PDesc *>(metric_all + 2 * WARPS * METRIC_DOUBLES); // [slot 2][warp]
    EndTma *ends_all = reinterpret_cast<EndTma *>(desc_all + 2 * WARPS);              // [slot 2][W,E][column, S diag, N diag]
    uint64_t *bars = reinterpret_cast<uint64_t *>(ends_all + 12);                      // full [warp][stage]
    uint64_t *cbars = bars + WARPS * STAGES;                                           // cons [warp][2]
    uint64_t *ibars = cbars + WARPS * 2;                                               // "item id published" [2]
    volatile int *wq = reinterpret_cast<volatile int *>(ibars + 2);                    // [0,1] published ids, [2,3] the CTA's first two items

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool w_end = warp == 0, e_end = warp == WARPS - 1, is_end = w_end || e_end;
    const int side = e_end ? 1 : 0;
    const int ntiles_all = (p.kq + TILE_SLABS - 1) / TILE_SLABS;
    const int ubase = ntiles_all / p.units, urem = ntiles_all % p.units; // unit u: tiles [u*ubase + min(u,urem), + ubase + (u < urem))
    const int nstrips = p.strip_count;
    const int nitems = nstrips * p.units;
    // Work items come from a global counter (dynamic): the items in flight are then always the ~444 most recently
    // fetched ones, i.e. a compact band of neighbouring strips of one unit whose shared lines are in L2.  (A static
    // round-robin assignment lets the CTAs drift apart over their ~1000 items: measured 4.1x DRAM reads, 26 ms.)
    if (threadIdx.x == 0) {
        if ((smem_u32(psm) & 1023u) != 0) __trap(); // the budget has no alignment slack
        wq[2] = atomicAdd(p.work_counter, 1);
        wq[3] = atomicAdd(p.work_counter, 1);
        mbar_init(smem_u32(&ibars[0]), 1);
        mbar_init(smem_u32(&ibars[1]), 1);
    }
    __syncthreads();
    int it = wq[2], nxt = wq[3];
    if (it >= nitems) return;

    const uint32_t stage0 = smem_u32(stages + (size_t)warp * STAGES * SB);
    const uint32_t bar0 = smem_u32(&bars[warp * STAGES]);
    const uint32_t cbar0 = smem_u32(&cbars[warp * 2]);
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(bar0 + s * 8, is_end ? 3u : 2u); // tile + S/N rows (+ strip-end boxes)
        mbar_init(cbar0, is_end ? 32u : 64u);                                       // every lane of every mate
        mbar_init(cbar0 + 8, is_end ? 32u : 64u);
        fence_barrier_init();
        fence_proxy_async();
    }

    // whole warp: decode item `item` (its StripReg word `raw` per lane) into descriptor slot sl
    auto decode = [&](int sl, int item, uint32_t raw) {
        auto W = [&](int i) { return __shfl_sync(0xffffffffu, raw, i); };
        const uint32_t first = W(29), sS = W(1 + warp), sN = W(5 + warp);
        const uint32_t c0S = (W(9) >> (8 * warp)) & 255u, c0N = (W(10) >> (8 * warp)) & 255u;
        const uint32_t cn = W(11 + warp), cp = W(15 + warp);
        const uint32_t w21 = W(21), w22 = W(22), w27 = W(27), w28 = W(28);
        const uint32_t esrc = W(19 + side), dsrc0 = W(23 + 2 * side), dsrc1 = W(24 + 2 * side);
        if (lane == 0) {
            const int u = item / nstrips;
            PDesc &D = desc_all[sl * WARPS + warp];
            D.le = (int32_t)first + warp;
            D.tile0 = u * ubase + min(u, urem);
            D.ntile = ubase + (u < urem ? 1 : 0);
            D.elemS = (int32_t)sS;
            D.elemN = (int32_t)sN;
            const uint32_t hS = c0S == (uint32_t)ROW_IN_HALO, hN = c0N == (uint32_t)ROW_IN_HALO;
            D.sel = (hS ? 0u : c0S) | ((hN ? 0u : c0N) << 8) | (hS << 16) | (hN << 17) | ((((w22 >> 16) & 255u) ? 1u : 0u) << 18);
            D.cn_pack = cn;
            D.cperm_pack = cp;
            uint32_t eb = 0;
            if (is_end) {
                EndTma *E = ends_all + (sl * 2 + side) * 3;
                const int kind = (w21 >> (8 * side)) & 255, row = (w21 >> (16 + 8 * side)) & 255;
                EndTma e{ET_NONE, 0, 0, 0, 0, 0, 0};
                if (kind == SRC_TMPCOL) e = {ET_COL, (uint8_t)(row ? 1 : 0), 0, 0, row ? 2 : 0, 0, 0};
                else if (kind == SRC_TMP) e = {ET_ROW, 0, 0, 0, 4 * row, 0, 0};
                else e = {ET_HALOROW, 0, 0, 0, 0, 0, 0};
                e.rev = (uint8_t)((w22 >> (8 * side)) & 255u);
                e.src = (int32_t)esrc;
                E[0] = e;
                eb += end_box_bytes(e.kind);
#pragma unroll
                for (int which = 1; which < 3; ++which) {
                    const int d = side * 2 + (which - 1);
                    const int32_t ds = (int32_t)(which == 1 ? dsrc0 : dsrc1);
                    EndTma f{ET_NONE, 0, 0, 0, 0, 0, 0};
                    if (ds >= 0) {
                        const int pt = (w28 >> (8 * d)) & 255, dk = (w27 >> (8 * d)) & 255;
                        if (dk == SRC_HALO) f = {ET_HALOPT, (uint8_t)(pt & 1), 0, 0, pt & 2, 0, 0};
                        else f = {ET_PT, (uint8_t)(pt & 1), 0, 0, pt & 2, pt >> 2, 0};
                        f.src = ds;
                    }
                    E[which] = f;
                    eb += end_box_bytes(f.kind);
                }
            }
            D.end_bytes = eb;
        }
        __syncwarp();
    };
    auto load_raw = [&](int item) { return __ldg(reinterpret_cast<const uint32_t *>(p.strip_reg + (item % nstrips)) + lane); };
    auto load_metric = [&](int sl) { // this warp's folded metric (512 B) of the item in slot sl, asynchronously
        const int le = desc_all[sl * WARPS + warp].le;
        cp_async16(smem_u32(metric_all + (sl * WARPS + warp) * METRIC_DOUBLES) + lane * 16, p.metric + (size_t)le * METRIC_DOUBLES + lane * 2, 16u);
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    // ---- lane 0's three producers; (sl, n) = descriptor slot and tile inside that item, s = ring stage / barrier of that tile
    auto issue_tile = [&](int sl, int n, int s) {
        const PDesc &D = desc_all[sl * WARPS + warp];
        const uint32_t bar = bar0 + s * 8;
        mbar_expect_tx(bar, (uint32_t)TILE_BYTES);
        tma_load_3d(stage0 + s * SB, &p.map_tmp, 0, (D.tile0 + n) * TILE_SLABS, D.le, bar);
    };
    auto issue_nbr = [&](int sl, int n, int s, int q) {
        const PDesc &D = desc_all[sl * WARPS + warp];
        const int kq0 = (D.tile0 + n) * TILE_SLABS;
        const uint32_t bar = bar0 + s * 8;
        const uint32_t rb = smem_u32(rowbuf) + (uint32_t)((q * STRIP_MAX + warp) * 2048);
        const uint32_t sel = D.sel;
        mbar_expect_tx(bar, 2048u);
        tma_load_3d(rb, (sel >> 16) & 1u ? &p.map_halo : &p.map_rows, (int)(sel & 255u), kq0, D.elemS, bar);
        tma_load_3d(rb + 1024, (sel >> 17) & 1u ? &p.map_halo : &p.map_rows, (int)((sel >> 8) & 255u), kq0, D.elemN, bar);
    };
    auto issue_end = [&](int sl, int n, int s) {
        const PDesc &D = desc_all[sl * WARPS + warp];
        const int kq0 = (D.tile0 + n) * TILE_SLABS;
        const uint32_t bar = bar0 + s * 8;
        mbar_expect_tx(bar, D.end_bytes);
#pragma unroll
        for (int which = 0; which < 3; ++which) {
            const EndTma e = ends_all[(sl * 2 + side) * 3 + which];
            const uint32_t dst = smem_u32(endbuf) + (uint32_t)(side * END_SIDE_BYTES + (which ? 1536 + 512 * which : 0));
            switch (e.kind) {
            case ET_COL:
#pragma unroll
                for (int j = 0; j < 4; ++j) tma_load_4d(dst + j * 512, &p.map_pt, e.c0, j, kq0, e.src, bar);
                break;
            case ET_ROW: tma_load_3d(dst, &p.map_rows, e.c0, kq0, e.src, bar); break;
            case ET_HALOROW: tma_load_3d(dst, &p.map_halo, 0, kq0, e.src, bar); break;
            case ET_PT: tma_load_4d(dst, &p.map_pt, e.c0, e.c1, kq0, e.src, bar); break;
            case ET_HALOPT: tma_load_3d(dst, &p.map_halo_pt, e.c0, kq0, e.src, bar); break;
            default: break;
            }
        }
    };

    decode(0, it, load_raw(it));
    load_metric(0);
    __syncthreads(); // the only CTA-wide barrier: every warp's mbarriers are initialised before a mate touches them
    if (lane == 0) { // fill the pipeline: tiles 0..2, rows 0..1, strip-end boxes 0 of the first item (every unit has >= 3 tiles)
#pragma unroll
        for (int n = 0; n < STAGES; ++n) issue_tile(0, n, n);
        issue_nbr(0, 0, 0, 0);
        if (p.nbr_ahead == 2) issue_nbr(0, 1, 1, 1);
        if (is_end) issue_end(0, 0, 0);
    }

    const int sw = lane & 7;         // SWIZZLE_128B
    const int swz = (lane >> 2) & 1; // SWIZZLE_32B
    const uint64_t pol_out = l2_policy_evict_first();
    int sl = 0, s = 0;
    uint32_t parity = 0, g = 0; // g: tiles this warp has done (rowbuf set g & 1, cons parity (g >> 1) & 1)

    uint32_t kitem = 0; // items this CTA has started
    while (true) {
        const bool has_next = nxt < nitems;
        uint32_t raw = 0;
        int pend = 0;
        if (has_next) {
            raw = load_raw(nxt); // consumed one tile before the look-ahead crosses into the next item
            if (threadIdx.x == 0) pend = atomicAdd(p.work_counter, 1); // the id of the item after the next; published below
        }
        const PDesc &D = desc_all[sl * WARPS + warp];
        const int ntile = D.ntile, tile0 = D.tile0, le = D.le;
        const uint32_t cn_pack = D.cn_pack, cperm_pack = D.cperm_pack;
        const bool std_corners = (D.sel >> 18) & 1u;
        const SmemMetric M{metric_all + (sl * WARPS + warp) * METRIC_DOUBLES};
        const EndTma *endtma = ends_all + sl * 6;

        auto read_end = [&](int sd, double (&col)[4], double &d0, double &d1) {
            const unsigned char *eb = endbuf + sd * END_SIDE_BYTES;
            const EndTma c = endtma[3 * sd];
            const int rm = c.rev ? 3 : 0;
            if (c.kind == ET_COL) {
                const unsigned char *b = eb + lane * 16 + c.hi * 8;
#pragma unroll
                for (int t = 0; t < 4; ++t) col[t] = *reinterpret_cast<const double *>(b + ((t ^ rm) << 9));
            } else {
                const unsigned char *b = eb + lane * 32;
                const int x = ((lane >> 2) & 1) ^ (rm >> 1), y = rm & 1;
#pragma unroll
                for (int t = 0; t < 4; ++t) col[t] = *reinterpret_cast<const double *>(b + (((t >> 1) ^ x) << 4) + (((t & 1) ^ y) << 3));
            }
            d0 = *reinterpret_cast<const double *>(eb + 2048 + lane * 16 + endtma[3 * sd + 1].hi * 8);
            d1 = *reinterpret_cast<const double *>(eb + 2560 + lane * 16 + endtma[3 * sd + 2].hi * 8);
        };

        for (int n = 0; n < ntile; ++n) {
            if (n == ntile - 2 && has_next) { // the look-ahead reaches the next item at the end of this iteration
                decode(sl ^ 1, nxt, raw);
                load_metric(sl ^ 1);
                if (threadIdx.x == 0) { // publish the id fetched at the top of this item (slot kitem & 1, read at the top of the next item)
                    wq[kitem & 1u] = pend;
                    mbar_arrive(smem_u32(&ibars[kitem & 1u]));
                }
            }
            const int q = (int)(g & 1u);
            const uint32_t qparity = (g >> 1) & 1u;
            const int kq0 = (tile0 + n) * TILE_SLABS;
            unsigned char *row = stages + ((size_t)warp * STAGES + s) * SB + lane * 128;
            const unsigned char *rb = rowbuf + (size_t)((q * STRIP_MAX + warp) * 2048) + lane * 32;
            double v[16], nb[4][4], dg[4];

            mbar_wait(bar0 + s * 8, parity);
            if (n == 0) { // this item's metric: staged by this warp's own cp.async one or two tiles ago
                asm volatile("cp.async.wait_all;" ::: "memory");
                __syncwarp();
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
                v[2 * c] = t.x;
                v[2 * c + 1] = t.y;
            }
            {
                const double2 a = *reinterpret_cast<const double2 *>(rb + (swz << 4)), b = *reinterpret_cast<const double2 *>(rb + ((1 ^ swz) << 4));
                nb[EDGE_S][0] = a.x; nb[EDGE_S][1] = a.y; nb[EDGE_S][2] = b.x; nb[EDGE_S][3] = b.y;
                const double2 c = *reinterpret_cast<const double2 *>(rb + 1024 + (swz << 4)), d = *reinterpret_cast<const double2 *>(rb + 1024 + ((1 ^ swz) << 4));
                nb[EDGE_N][0] = c.x; nb[EDGE_N][1] = c.y; nb[EDGE_N][2] = d.x; nb[EDGE_N][3] = d.y;
            }
            if (w_end) {
                read_end(0, nb[EDGE_W], dg[0], dg[2]);
            } else {
                mbar_wait(bar0 - STAGES * 8 + s * 8, parity);
                const unsigned char *mrow = row - STAGES * SB;
#pragma unroll
                for (int t = 0; t < 4; ++t) nb[EDGE_W][t] = *reinterpret_cast<const double *>(mrow + (((2 * t + 1) ^ sw) << 4) + 8);
                dg[0] = *reinterpret_cast<const double *>(rb - 2048 + ((1 ^ swz) << 4) + 8);
                dg[2] = *reinterpret_cast<const double *>(rb - 2048 + 1024 + ((1 ^ swz) << 4) + 8);
            }
            if (e_end) {
                read_end(1, nb[EDGE_E], dg[1], dg[3]);
            } else {
                mbar_wait(bar0 + STAGES * 8 + s * 8, parity);
                const unsigned char *mrow = row + STAGES * SB;
#pragma unroll
                for (int t = 0; t < 4; ++t) nb[EDGE_E][t] = *reinterpret_cast<const double *>(mrow + (((2 * t) ^ sw) << 4));
                dg[1] = *reinterpret_cast<const double *>(rb + 2048 + (swz << 4));
                dg[3] = *reinterpret_cast<const double *>(rb + 2048 + 1024 + (swz << 4));
            }

            // ---- DSS (edgeVunpack), corners in ascending global point order
            v[1] += nb[EDGE_S][1]; v[2] += nb[EDGE_S][2];
            v[13] += nb[EDGE_N][1]; v[14] += nb[EDGE_N][2];
            v[4] += nb[EDGE_W][1]; v[8] += nb[EDGE_W][2];
            v[7] += nb[EDGE_E][1]; v[11] += nb[EDGE_E][2];
            if (std_corners) {
                v[0] = ((dg[0] + nb[EDGE_S][0]) + nb[EDGE_W][0]) + v[0];
                v[3] = ((nb[EDGE_S][3] + dg[1]) + v[3]) + nb[EDGE_E][0];
                v[12] = ((nb[EDGE_W][3] + v[12]) + dg[2]) + nb[EDGE_N][0];
                v[15] = ((v[15] + nb[EDGE_E][3]) + nb[EDGE_N][3]) + dg[3];
            } else {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int pt = c == 0 ? 0 : (c == 1 ? 3 : (c == 2 ? 12 : 15));
                    const double cv0 = v[pt];
                    const double cv1 = nb[c < 2 ? EDGE_S : EDGE_N][(c & 1) ? 3 : 0];
                    const double cv2 = nb[(c & 1) ? EDGE_E : EDGE_W][c < 2 ? 0 : 3];
                    const double cv3 = dg[c];
                    const int perm = (cperm_pack >> (8 * c)) & 255, cn = (cn_pack >> (8 * c)) & 255;
                    auto pick = [&](int k) {
                        const int w = (perm >> (2 * k)) & 3;
                        return w == 0 ? cv0 : (w == 1 ? cv1 : (w == 2 ? cv2 : cv3));
                    };
                    double sum = pick(0);
                    if (cn > 1) sum = sum + pick(1);
                    if (cn > 2) sum = sum + pick(2);
                    if (cn > 3) sum = sum + pick(3);
                    v[pt] = sum;
                }
            }
            // the mates' values have been consumed: tell them (the operands tie the arrival to the sums)
            if (!w_end) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(cbar0 - 16 + q * 8), "d"(v[0]), "d"(v[12]), "d"(v[4]), "d"(v[8]) : "memory");
            if (!e_end) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(cbar0 + 16 + q * 8), "d"(v[3]), "d"(v[15]), "d"(v[7]), "d"(v[11]) : "memory");
            const int s1 = s + 1 == STAGES ? 0 : s + 1, s2 = s == 0 ? STAGES - 1 : s - 1; // stages of tiles g+1 and g+2 (= g-1)
            if (is_end) { // endbuf consumed by every lane: the next tile's boxes travel while this tile's Laplacian runs
                asm volatile("" ::"d"(v[0]), "d"(v[12]), "d"(v[4]), "d"(v[8]), "d"(v[3]), "d"(v[15]), "d"(v[7]), "d"(v[11]) : "memory");
                __syncwarp();
                if (lane == 0) {
                    if (n + 1 < ntile) issue_end(sl, n + 1, s1);
                    else if (has_next) issue_end(sl ^ 1, 0, s1);
                }
            }

            if (KIND != SK_DSS) laplace_slab(v, M);
            mbar_wait(cbar0 + q * 8, qparity); // both mates have read this tile's columns and this rowbuf set
#pragma unroll
            for (int c = 0; c < 8; ++c) *reinterpret_cast<double2 *>(row + ((c ^ sw) << 4)) = make_double2(v[2 * c], v[2 * c + 1]);
            fence_proxy_async();
            __syncwarp();

            if (lane == 0) {
                tma_store_3d_hint(&p.map_out, 0, kq0, le, stage0 + s * SB, pol_out);
                bulk_commit();
                // tile g+2: into the stage tile g-1 used (once its store has read it) and into rowbuf set q
                int sl2 = sl, n2 = n + 2;
                bool have2 = true;
                if (n2 >= ntile) { sl2 = sl ^ 1; n2 -= ntile; have2 = has_next; }
                if (have2) {
                    if (g >= 1) {
                        bulk_wait_read<1>();
                        issue_tile(sl2, n2, s2);
                    }
                    if (p.nbr_ahead == 2) issue_nbr(sl2, n2, s2, q);
                }
                if (p.nbr_ahead != 2) { // rows one tile ahead: the owner of those lines has usually asked for them by then
                    if (n + 1 < ntile) issue_nbr(sl, n + 1, s1, q ^ 1);
                    else if (has_next) issue_nbr(sl ^ 1, 0, s1, q ^ 1);
                }
            }
            s = s1;
            if (s == 0) parity ^= 1u;
            ++g;
        }
        if (!has_next) break;
        it = nxt;
        mbar_wait(smem_u32(&ibars[kitem & 1u]), (kitem >> 1) & 1u);
        nxt = wq[kitem & 1u];
        ++kitem;
        sl ^= 1;
    }
    if (lane == 0) bulk_wait_read<0>();
}

// ------------------------------------------------------------------------------------------
// Warp-strip form of the regular strip sweep (round 2): ONE WARP owns all four members of a strip.
// Lane l = (member l >> 3, slab l & 7): a warp's tile is 8 consecutive (k,q) slabs of 4 consecutive elements, one
// 4 KB TMA box {16, 8 slabs, 4 elements}; row r of the staged tile is lane r's slab, so the swizzled 128-byte reads are
// those of the element-local sweeps.  The W/E edge exchange between strip mates and the diagonal corner values are
// warp shuffles by 8 lanes -- no shared-memory column reads, no mate / cons mbarriers, no in-place hazard: the four
// warps of a CTA are completely independent pipelines that only share the strip's staged metric.  (In the one-warp-
// per-member form the largest stalls were the waits on the mates' barriers: 16 % of the samples.)
//   * S/N neighbour rows of the four members: one {4, 8, 4} box each (the neighbours of consecutive members of a
//     face-interior strip are consecutive elements; the host sends every other strip to sweep2_strip_kernel);
//   * strip ends: only the 16 lanes of the two end members need the outside neighbour's facing column and the two
//     diagonal values (6 doubles per lane and tile).  They fetch them themselves with 8-byte cp.async one tile ahead
//     into private slots -- every source kind and orientation is folded into one base/stride record per strip end.
//     (Staging them by TMA, as the one-warp-per-member form does, cost ~150 of 864 instructions per tile here: six
//     small boxes issued by lane 0 and a kind switch on both sides.);
//   * a CTA = 4 warps on the same strip, warp w takes slabs [8w, 8w+8) of every 32-slab tile of the unit, so the
//     CTA still streams 4 KB contiguous per element and tile.
// ------------------------------------------------------------------------------------------
// NM members per warp (4, 8 or 16 = 1, 2 or 4 consecutive strips), 32 / NM slabs of each per warp tile: the strip ends'
// share of the work falls with NM (16, 8 or 4 end lanes per warp)
constexpr int WS_ROWBUF_BYTES = 2 * 4 * 2 * 1024;             // [set][warp][S,N][32 rows x 32 B]
template <int NM> constexpr int ws_endbuf_bytes() { return 4 * 2 * 2 * (32 / NM) * 48; } // [warp][set][end lane][4 edge values, 2 diagonals]
constexpr int WS_METRIC_STRIDE = METRIC_DOUBLES + 2;          // doubles; +16 B per member: the four members' reads hit different banks
// Source of one strip end, decoded once per CTA: the outside neighbour's four values along the facing edge in THIS
// element's orientation are at col + slab * slab_bytes + t * step (a column or a row of a local slab, or a received halo
// row, either direction), the two diagonal corner values at d0 / d1 + slab * d?_slab_bytes (nullptr = none, cube corner).
struct WsEnd {
    const unsigned char *col, *d0, *d1;
    int32_t step, slab_bytes, d0_slab_bytes, d1_slab_bytes;
    int32_t pad[2];
};
static_assert(sizeof(WsEnd) == 48, "WsEnd layout");
// ROWSETS = 2: S/N rows requested two tiles ahead (double buffered); 1: one set, requested right after the previous tile's rows
// have been consumed -- with a 2-stage tile ring that is the 50 KB / 128-register shape that puts FOUR CTAs on an SM
template <int STAGES, int ROWSETS, int NM, bool POST>
constexpr int wstrip_smem_bytes()
{
    return 1024 + 4 * STAGES * TILE_BYTES + ROWSETS * (WS_ROWBUF_BYTES / 2) + ws_endbuf_bytes<NM>() + 4 * (STAGES + 1) * 8 + NM * WS_METRIC_STRIDE * 8 + 2 * 48 +
           (POST ? NM * 34 * 8 : 0);
}
template <int D> __device__ __forceinline__ double shfl_up_d(double x) { return __shfl_up_sync(0xffffffffu, x, D); }
template <int D> __device__ __forceinline__ double shfl_down_d(double x) { return __shfl_down_sync(0xffffffffu, x, D); }

template <int STAGES, int KIND, int MINB, int NM>
__global__ void __launch_bounds__(128, MINB) sweep2_wstrip_kernel(const __grid_constant__ SweepParams p)
{
    constexpr int SB = TILE_BYTES, WARPS = 4, ROWSETS = MINB >= 4 ? 1 : 2;
    // S/N rows: two {4, SL, NM} TMA boxes per tile issued by lane 0.  The alternative kept below (ROWS_BY_TMA = false: every lane
    // fetches the 64 bytes of its own slab's S and N rows with four 16-byte cp.async in the group of the strip-end values, no
    // lane-0 code, one mbarrier arrival per tile) is correct but SLOWER: 13.6-14.0 ms against 11.4-11.8 ms at ne120 with 8
    // members (profiles/r2w_*): a warp-wide 16-byte gather touches 32 different lines, and the L1/LSU path pays for each.
    constexpr bool ROWS_BY_TMA = true;
    constexpr int SL = 32 / NM, NG = NM / 4;   // slabs of a member per warp tile; strips (StripReg records) per warp
    constexpr int ENDSET = 2 * SL * 48, ENDWARP = 2 * ENDSET;
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char *stages = smem;                                               // [warp][stage][4 members x 8 slabs x 128 B], SWIZZLE_128B
    unsigned char *rowbuf = stages + WARPS * STAGES * SB;                       // [set][warp][S,N][32 x 32 B], SWIZZLE_32B
    unsigned char *endbuf = rowbuf + ROWSETS * (WS_ROWBUF_BYTES / 2);                           // [warp][set][end lane][48 B]: cp.async slots of the end members
    uint64_t *bars = reinterpret_cast<uint64_t *>(endbuf + WARPS * ENDWARP);    // full [warp][stage], qdp_full [warp]
    double *metric_s = reinterpret_cast<double *>(bars + WARPS * (STAGES + 1)); // [member][WS_METRIC_STRIDE]
    WsEnd *wsend = reinterpret_cast<WsEnd *>(metric_s + NM * WS_METRIC_STRIDE); // [W end, E end]
    double *hv_s = reinterpret_cast<double *>(wsend + 2);                       // SK_POST: [member][spheremp 16 | rspheremp 16 | pad 2]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int member = lane / SL, j = lane % SL;
    const StripReg *regs = p.strip_reg + (size_t)blockIdx.x * NG; // NG consecutive strips: members 4g .. 4g+3 are strip g's
    const StripReg &rg = regs[0], &rgl = regs[NG - 1], &rgm = regs[member >> 2];
    const int first = rg.first;
    const int ntiles = (p.kq + TILE_SLABS - 1) / TILE_SLABS;
    const int t0 = blockIdx.y * p.tiles_per_unit;
    const int t1 = min(t0 + p.tiles_per_unit, ntiles);
    // this warp's tiles: slabs [t0 * 32 + n * 4 SL + SL warp, + SL) of the NM members (a CTA step covers 4 SL consecutive slabs)
    int my_n = (t1 - t0) * (8 / SL);
    while (my_n > 0 && t0 * TILE_SLABS + (my_n - 1) * 4 * SL + warp * SL >= p.kq) --my_n;
    const uint32_t stage0 = smem_u32(stages + (size_t)warp * STAGES * SB);
    const uint32_t bar0 = smem_u32(&bars[warp * STAGES]);
    const uint32_t qbar = smem_u32(&bars[WARPS * STAGES + warp]);
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(bar0 + s * 8, ROWS_BY_TMA ? 2u : 1u); // tile (+ S/N rows when TMA stages them)
        mbar_init(qbar, 1u);
        fence_barrier_init();
        fence_proxy_async();
        const int pre = my_n < STAGES ? my_n : STAGES;
        for (int n = 0; n < pre; ++n) {
            mbar_expect_tx(bar0 + n * 8, (uint32_t)TILE_BYTES);
            tma_load_3d(stage0 + n * SB, &p.map4_tmp, 0, t0 * TILE_SLABS + n * 4 * SL + warp * SL, first, bar0 + n * 8);
        }
    }
    for (int t = threadIdx.x; t < NM * 32; t += WARPS * 32) { // the members' folded metric (16 bytes per item), shared by the four warps
        const int m = t >> 5, c = (t & 31) * 2;
        *reinterpret_cast<double2 *>(metric_s + m * WS_METRIC_STRIDE + c) = __ldg(reinterpret_cast<const double2 *>(p.metric + (size_t)(first + m) * METRIC_DOUBLES + c));
        if (KIND == SK_POST) {
            const int l = t & 31;
            hv_s[m * 34 + l] = l < 16 ? p.elem[(size_t)(first + m) * ELEM_DOUBLES + 64 + l] : p.rsph[(size_t)(first + m) * 16 + l - 16];
        }
    }
    if (threadIdx.x < 2) { // the two strip ends: fold source kind, row/column and orientation into base + strides
        const int side = threadIdx.x;
        const size_t kq = (size_t)p.kq;
        const unsigned char *tmp = reinterpret_cast<const unsigned char *>(p.tmp), *halo = reinterpret_cast<const unsigned char *>(p.halo);
        WsEnd e{nullptr, nullptr, nullptr, 0, 0, 0, 0, {0, 0}};
        const StripReg &re = side ? rgl : rg; // W end: the first strip's record; E end: the last strip's
        const int kind = re.end_kind[side], row = re.end_row[side], rev = re.end_rev[side];
        const size_t src = (size_t)re.end_src[side];
        if (kind == SRC_TMPCOL) { // the neighbour's i = 1 (row 0) or i = np (row 1) column: points i + 4 t
            e.col = tmp + src * kq * 128 + ((row ? 3 : 0) + (rev ? 12 : 0)) * 8;
            e.step = rev ? -32 : 32;
            e.slab_bytes = 128;
        } else if (kind == SRC_TMP) { // a natural j-row of the neighbour's slab
            e.col = tmp + src * kq * 128 + (4 * row + (rev ? 3 : 0)) * 8;
            e.step = rev ? -8 : 8;
            e.slab_bytes = 128;
        } else { // a received halo row
            e.col = halo + src * kq * 32 + (rev ? 24 : 0);
            e.step = rev ? -8 : 8;
            e.slab_bytes = 32;
        }
#pragma unroll
        for (int which = 0; which < 2; ++which) {
            const int d = side * 2 + which;
            const unsigned char *ptr = nullptr;
            int sb = 0;
            if (re.diag_src[d] >= 0) {
                const size_t ds = (size_t)re.diag_src[d];
                if (re.diag_kind[d] == SRC_HALO) { ptr = halo + ds * kq * 32 + re.diag_pt[d] * 8; sb = 32; }
                else { ptr = tmp + ds * kq * 128 + re.diag_pt[d] * 8; sb = 128; }
            }
            if (which == 0) { e.d0 = ptr; e.d0_slab_bytes = sb; }
            else { e.d1 = ptr; e.d1_slab_bytes = sb; }
        }
        if (p.debug_flags & 2) {
            e.col = tmp + (size_t)first * kq * 128;
            e.step = 8;
            e.slab_bytes = 128;
            e.d0 = e.d1 = e.col;
            e.d0_slab_bytes = e.d1_slab_bytes = 128;
        }
        wsend[side] = e;
    }
    __syncthreads(); // the only CTA-wide barrier
    if (my_n <= 0) return;

    const int swz = (lane >> 2) & 1, sw = lane & 7;
    // (p.debug_flags, timing experiments only, results are wrong: 1 = the S/N rows are read from the strip's own slabs,
    //  2 = the strip-end values are read from the strip's own first slab: no line of another CTA is touched)
    const int elemS = (p.debug_flags & 1) ? first : rg.s_src[0], elemN = (p.debug_flags & 1) ? first : rg.n_src[0];
    const int rawS = (p.debug_flags & 1) ? 0 : rg.s_c0[0], rawN = (p.debug_flags & 1) ? 12 : rg.n_c0[0];
    const CUtensorMap *mapS = rawS == ROW_IN_HALO ? &p.map4_halo : &p.map4_rows, *mapN = rawN == ROW_IN_HALO ? &p.map4_halo : &p.map4_rows;
    const int c0S = rawS == ROW_IN_HALO ? 0 : rawS, c0N = rawN == ROW_IN_HALO ? 0 : rawN;
    bool std_corners = true;
#pragma unroll
    for (int g = 0; g < NG; ++g) std_corners = std_corners && regs[g].all_std != 0;
    uint32_t cn_pack = 0, cperm_pack = 0; // this lane's member: four corner counts / summation orders
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        cn_pack |= (uint32_t)rgm.corner_n[member & 3][c] << (8 * c);
        cperm_pack |= (uint32_t)rgm.corner_perm[member & 3][c] << (8 * c);
    }
    const uint64_t pol_out = l2_policy_evict_first();
    // end-member lanes: slot of this lane in the warp's two cp.async sets, and the strip end it serves
    const bool end_lane = member == 0 || member == NM - 1;
    const int eside = member == NM - 1 ? 1 : 0;
    const uint32_t eslot0 = smem_u32(endbuf) + (uint32_t)(warp * ENDWARP + (eside * SL + j) * 48);
    const unsigned char *eslot0_p = endbuf + warp * ENDWARP + (eside * SL + j) * 48;
    auto kq_of = [&](int n) { return t0 * TILE_SLABS + n * 4 * SL + warp * SL; };
    auto issue_tile = [&](int n, int s) {
        mbar_expect_tx(bar0 + s * 8, (uint32_t)TILE_BYTES);
        tma_load_3d(stage0 + s * SB, &p.map4_tmp, 0, kq_of(n), first, bar0 + s * 8);
    };
    auto issue_nbr = [&](int n) {
        const uint32_t nbar = bar0 + (n % STAGES) * 8;
        const uint32_t rb = smem_u32(rowbuf) + (uint32_t)(((ROWSETS == 2 ? (n & 1) * WARPS : 0) + warp) * 2048);
        mbar_expect_tx(nbar, 2048u);
        tma_load_3d(rb, mapS, c0S, kq_of(n), elemS, nbar);
        tma_load_3d(rb + 1024, mapN, c0N, kq_of(n), elemN, nbar);
    };
    // end-member lanes: fetch the strip-end values of tile n into set n & 1 (six 8-byte cp.async, one group)
    const unsigned char *rowS_src = rawS == ROW_IN_HALO ? reinterpret_cast<const unsigned char *>(p.halo) + (size_t)(elemS + member) * p.kq * 32
                                                        : reinterpret_cast<const unsigned char *>(p.tmp) + (size_t)(elemS + member) * p.kq * 128 + c0S * 8;
    const unsigned char *rowN_src = rawN == ROW_IN_HALO ? reinterpret_cast<const unsigned char *>(p.halo) + (size_t)(elemN + member) * p.kq * 32
                                                        : reinterpret_cast<const unsigned char *>(p.tmp) + (size_t)(elemN + member) * p.kq * 128 + c0N * 8;
    const int rowS_stride = rawS == ROW_IN_HALO ? 32 : 128, rowN_stride = rawN == ROW_IN_HALO ? 32 : 128;
    auto fetch_end = [&](int n) {
        if (!ROWS_BY_TMA) { // this lane's S and N rows of tile n, into the SWIZZLE_32B positions the TMA form uses
            const int slab = kq_of(n) + j;
            const uint32_t ok = slab < p.kq ? 16u : 0u;
            const size_t sl = ok ? (size_t)slab : 0;
            const uint32_t rb = smem_u32(rowbuf) + (uint32_t)(((n & 1) * WARPS + warp) * 2048) + lane * 32;
            const unsigned char *a = rowS_src + sl * rowS_stride, *b = rowN_src + sl * rowN_stride;
            cp_async16(rb + (swz << 4), a, ok);
            cp_async16(rb + ((1 ^ swz) << 4), a + 16, ok);
            cp_async16(rb + 1024 + (swz << 4), b, ok);
            cp_async16(rb + 1024 + ((1 ^ swz) << 4), b + 16, ok);
        }
        if (end_lane) {
            const WsEnd e = wsend[eside];
            const int slab = kq_of(n) + j;
            const uint32_t ok = slab < p.kq ? 8u : 0u; // slabs past a ragged end: zero fill, nothing is read
            const size_t sl = ok ? (size_t)slab : 0;
            const uint32_t dst = eslot0 + (uint32_t)((n & 1) * ENDSET);
            const unsigned char *c = e.col + sl * (size_t)e.slab_bytes;
#pragma unroll
            for (int t = 0; t < 4; ++t) cp_async8(dst + t * 8, c + t * e.step, ok);
            if (e.d0) cp_async8(dst + 32, e.d0 + sl * (size_t)e.d0_slab_bytes, ok);
            if (e.d1) cp_async8(dst + 40, e.d1 + sl * (size_t)e.d1_slab_bytes, ok);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    if (ROWS_BY_TMA && lane == 0) {
        issue_nbr(0);
        if (ROWSETS == 2 && my_n > 1) issue_nbr(1);
    }
    fetch_end(0);
    const SmemMetric M{metric_s + member * WS_METRIC_STRIDE};

    for (int n = 0; n < my_n; ++n) {
        const int s = n % STAGES, q = n & 1;
        const int kq0 = kq_of(n);
        const uint32_t parity = (uint32_t)((n / STAGES) & 1);
        unsigned char *row = stages + ((size_t)warp * STAGES + s) * SB + lane * 128;
        const unsigned char *rb = rowbuf + (size_t)(((ROWSETS == 2 ? q * WARPS : 0) + warp) * 2048) + lane * 32;
        double v[16], nbS[4], nbN[4], nbW[4], nbE[4], dg[4];

        mbar_wait(bar0 + s * 8, parity);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
            v[2 * c] = t.x;
            v[2 * c + 1] = t.y;
        }
        asm volatile("cp.async.wait_all;" ::: "memory"); // this lane's S/N rows and strip-end values of this tile (fetched one tile ago)
        {
            const double2 a = *reinterpret_cast<const double2 *>(rb + (swz << 4)), b = *reinterpret_cast<const double2 *>(rb + ((1 ^ swz) << 4));
            nbS[0] = a.x; nbS[1] = a.y; nbS[2] = b.x; nbS[3] = b.y;
            const double2 c = *reinterpret_cast<const double2 *>(rb + 1024 + (swz << 4)), d = *reinterpret_cast<const double2 *>(rb + 1024 + ((1 ^ swz) << 4));
            nbN[0] = c.x; nbN[1] = c.y; nbN[2] = d.x; nbN[3] = d.y;
        }
        if (ROWSETS == 1) { // the single row set has been read by every lane: it can take the next tile's rows
            asm volatile("" ::"d"(nbS[0]), "d"(nbS[2]), "d"(nbN[0]), "d"(nbN[2]) : "memory");
            __syncwarp();
            if (lane == 0 && n + 1 < my_n) {
                issue_nbr(n + 1);
                if (STAGES == 2 && n >= 1) { // 2-stage ring: tile n+1 goes where tile n-1 was, once that tile's store has read it
                    bulk_wait_read<0>();
                    issue_tile(n + 1, (n + 1) % STAGES);
                }
            }
        }
        // ---- W/E mates are SL lanes away: their facing columns and the facing ends of their S/N neighbour rows by shuffle
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            nbW[t] = shfl_up_d<SL>(v[3 + 4 * t]);   // west mate's E column (points 3,7,11,15)
            nbE[t] = shfl_down_d<SL>(v[4 * t]);     // east mate's W column (points 0,4,8,12)
        }
        dg[0] = shfl_up_d<SL>(nbS[3]);   // SW: east end of the west mate's S row
        dg[2] = shfl_up_d<SL>(nbN[3]);   // NW
        dg[1] = shfl_down_d<SL>(nbS[0]); // SE: west end of the east mate's S row
        dg[3] = shfl_down_d<SL>(nbN[0]); // NE
        // strip ends: the outside neighbour's values, fetched by this lane's own cp.async one tile ago
        if (end_lane) {
            const double2 *es = reinterpret_cast<const double2 *>(eslot0_p + q * ENDSET);
            const double2 c01 = es[0], c23 = es[1], dd = es[2];
            if (member == 0) { nbW[0] = c01.x; nbW[1] = c01.y; nbW[2] = c23.x; nbW[3] = c23.y; dg[0] = dd.x; dg[2] = dd.y; }
            else { nbE[0] = c01.x; nbE[1] = c01.y; nbE[2] = c23.x; nbE[3] = c23.y; dg[1] = dd.x; dg[3] = dd.y; }
        }
        if (n + 1 < my_n) fetch_end(n + 1);

        // ---- DSS (edgeVunpack): mid-edge points, then corners in ascending global point order
        v[1] += nbS[1]; v[2] += nbS[2];
        v[13] += nbN[1]; v[14] += nbN[2];
        v[4] += nbW[1]; v[8] += nbW[2];
        v[7] += nbE[1]; v[11] += nbE[2];
        if (std_corners) {
            v[0] = ((dg[0] + nbS[0]) + nbW[0]) + v[0];   // PERM_SW
            v[3] = ((nbS[3] + dg[1]) + v[3]) + nbE[0];   // PERM_SE
            v[12] = ((nbW[3] + v[12]) + dg[2]) + nbN[0]; // PERM_NW
            v[15] = ((v[15] + nbE[3]) + nbN[3]) + dg[3]; // PERM_NE
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int pt = c == 0 ? 0 : (c == 1 ? 3 : (c == 2 ? 12 : 15));
                const double cv0 = v[pt];
                const double cv1 = (c < 2 ? nbS : nbN)[(c & 1) ? 3 : 0];
                const double cv2 = ((c & 1) ? nbE : nbW)[c < 2 ? 0 : 3];
                const double cv3 = dg[c];
                const int perm = (cperm_pack >> (8 * c)) & 255, cn = (cn_pack >> (8 * c)) & 255;
                auto pick = [&](int k) {
                    const int w = (perm >> (2 * k)) & 3;
                    return w == 0 ? cv0 : (w == 1 ? cv1 : (w == 2 ? cv2 : cv3));
                };
                double sum = pick(0);
                if (cn > 1) sum = sum + pick(1);
                if (cn > 2) sum = sum + pick(2);
                if (cn > 3) sum = sum + pick(3);
                v[pt] = sum;
            }
        }
        if (KIND == SK_POST) { // every lane has its tile row in registers: the stage can take the slab of Qdp while the Laplacian runs
            asm volatile("" ::"d"(v[0]), "d"(v[3]), "d"(v[5]), "d"(v[6]), "d"(v[9]), "d"(v[10]), "d"(v[12]), "d"(v[15]) : "memory");
            __syncwarp();
            if (lane == 0) {
                mbar_expect_tx(qbar, (uint32_t)TILE_BYTES);
                tma_load_3d(stage0 + s * SB, &p.map4_q, 0, kq0, first, qbar);
            }
        }

        if (KIND != SK_DSS) laplace_slab(v, M);

        if (KIND == SK_POST) {
            mbar_wait(qbar, (uint32_t)(n & 1));
            double a[16];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
                a[2 * c] = t.x;
                a[2 * c + 1] = t.y;
            }
            hypervis_update_slab(v, a, hv_s + member * 34, hv_s + member * 34 + 16, p.hv_coef, p.hv_limiter);
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<double2 *>(row + ((c ^ sw) << 4)) = make_double2(v[2 * c], v[2 * c + 1]);
        fence_proxy_async();
        __syncwarp();

        if (lane == 0) {
            tma_store_3d_hint(&p.map4_out, 0, kq0, first, stage0 + s * SB, pol_out);
            bulk_commit();
            if (STAGES > 2 && n >= 1 && n - 1 + STAGES < my_n) { // refill the stage tile n-1 used, once its store has read it
                bulk_wait_read<1>();
                issue_tile(n - 1 + STAGES, (n - 1) % STAGES);
            }
            if (ROWS_BY_TMA && ROWSETS == 2 && n + 2 < my_n) issue_nbr(n + 2); // rowbuf set q is free: every lane consumed it before the __syncwarp above
        }
    }
    if (lane == 0) bulk_wait_read<0>();
}

// kernel configurations: (STAGES, WARPS, MINB)
#define BWK_CFG_LAPLACE 3, 4, 3
#define BWK_CFG_SWEEP1 3, 4, 3
#define BWK_CFG_SWEEP2 3, 4, 3

static int g_sweep2_tpu = 0; // 0: caller's hint / default; > 0: forced (BWK_SWEEP2_TPU)
void set_sweep2_tiles_per_unit(int tpu) { g_sweep2_tpu = tpu < 0 ? 0 : tpu; }
int sweep2_tiles_per_unit_forced() { return g_sweep2_tpu; }

SweepGeometry choose_geometry(int kq, int64_t count, int mode, int tpu_hint)
{
    const int warps = 4;
    const int ntiles = (kq + TILE_SLABS - 1) / TILE_SLABS;
    SweepGeometry g;
    if (mode == MODE_SWEEP2 || mode == MODE_DSS || mode == MODE_SWEEP2_POST) {
        // count = strips; every warp walks all tiles of the unit.  Long units amortise the per-CTA
        // set-up (descriptor, metric, pipeline fill); p.tiles_per_unit_hint overrides for tuning.
        int64_t tpu = g_sweep2_tpu > 0 ? g_sweep2_tpu : (tpu_hint > 0 ? tpu_hint : 15);
        const int64_t target_ctas = 148 * 8;
        if ((ntiles + tpu - 1) / tpu * count < target_ctas) tpu = std::max<int64_t>(1, (int64_t)ntiles * count / target_ctas);
        if (tpu > ntiles) tpu = ntiles;
        g.units = (int)((ntiles + tpu - 1) / tpu);
        g.tiles_per_unit = (int)((ntiles + g.units - 1) / g.units);
        return g;
    }
    // count = elements; the 4 warps of a CTA interleave the unit's tiles.  Enough CTAs for ~8 per
    // SM, but at most ~6 tiles per warp (short CTAs balance better)
    const int64_t target_ctas = 148 * 8;
    int64_t tpu = ((int64_t)ntiles * count + target_ctas - 1) / target_ctas;
    if (tpu < warps) tpu = warps;
    if (tpu > 6 * warps) tpu = 6 * warps;
    if (tpu > ntiles) tpu = ntiles;
    g.units = (int)((ntiles + tpu - 1) / tpu);
    g.tiles_per_unit = (int)((ntiles + g.units - 1) / g.units); // rebalance
    return g;
}

template <int MODE, int STAGES, int WARPS, int MINB>
static cudaError_t launch_sweep_t(const SweepParams &p, cudaStream_t st)
{
    const int smem = sweep_smem_bytes<STAGES, WARPS>() + (MODE == MODE_SWEEP1_PRE ? p.nlev * 128 : 0);
    static int configured = 0;
    if (configured < smem) {
        cudaError_t e = cudaFuncSetAttribute(sweep_kernel<MODE, STAGES, WARPS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    if (p.elem_count <= 0 || p.units <= 0) return cudaSuccess;
    dim3 grid((unsigned)p.elem_count, (unsigned)p.units);
    sweep_kernel<MODE, STAGES, WARPS, MINB><<<grid, WARPS * 32, smem, st>>>(p);
    return cudaGetLastError();
}

template <int STAGES, int WARPS, int MINB, int KIND>
static cudaError_t launch_strip_t(const SweepParams &p, cudaStream_t st)
{
    static_assert(WARPS == STRIP_MAX, "one warp per strip element");
    constexpr int smem = strip_smem_bytes<STAGES, WARPS>();
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(sweep2_strip_kernel<STAGES, WARPS, MINB, KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (p.strip_count <= 0 || p.units <= 0) return cudaSuccess;
    dim3 grid((unsigned)p.strip_count, (unsigned)p.units);
    sweep2_strip_kernel<STAGES, WARPS, MINB, KIND><<<grid, WARPS * 32, smem, st>>>(p);
    return cudaGetLastError();
}

template <int STAGES, int WARPS, int KIND>
static cudaError_t launch_persist_t(const SweepParams &p, cudaStream_t st)
{
    constexpr int smem = persist_smem_bytes<STAGES, WARPS>();
    static_assert(3 * (smem + 1024) <= 228 * 1024, "three CTAs per SM");
    static int ctas = 0;
    if (!ctas) {
        cudaError_t e = cudaFuncSetAttribute(sweep2_persist_kernel<STAGES, WARPS, KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sweep2_persist_kernel<STAGES, WARPS, KIND>, WARPS * 32, smem);
        if (e != cudaSuccess) return e;
        ctas = sms * (per_sm > 0 ? per_sm : 1);
    }
    if (p.strip_count <= 0 || p.units <= 0) return cudaSuccess;
    const long long nitems = (long long)p.strip_count * p.units;
    const int grid = (int)(nitems < ctas ? nitems : ctas);
    sweep2_persist_kernel<STAGES, WARPS, KIND><<<grid, WARPS * 32, smem, st>>>(p);
    return cudaGetLastError();
}

// persistent strip sweep over a list of FULL REGULAR strips (p.strip_reg / p.strip_count); p.units = number of units of the
// (k,q) axis, each of at least 3 tiles.  mode: MODE_SWEEP2 or MODE_DSS
cudaError_t launch_sweep2_persist(int mode, const SweepParams &p, cudaStream_t st)
{
    const int ntiles = (p.kq + TILE_SLABS - 1) / TILE_SLABS;
    if (p.units < 1 || ntiles / p.units < 3) return cudaErrorInvalidValue;
    if (mode == MODE_SWEEP2) return launch_persist_t<3, 4, SK_LAP>(p, st);
    if (mode == MODE_DSS) return launch_persist_t<3, 4, SK_DSS>(p, st);
    return cudaErrorInvalidValue;
}

template <int STAGES, int KIND, int MINB, int NM>
static cudaError_t launch_wstrip_t(const SweepParams &p, cudaStream_t st)
{
    constexpr int smem = wstrip_smem_bytes<STAGES, (MINB >= 4 ? 1 : 2), NM, KIND == SK_POST>();
    static_assert(MINB * (smem + 1024) <= 228 * 1024, "MINB CTAs per SM");
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(sweep2_wstrip_kernel<STAGES, KIND, MINB, NM>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (p.strip_count <= 0 || p.units <= 0) return cudaSuccess;
    dim3 grid((unsigned)p.strip_count, (unsigned)p.units);
    sweep2_wstrip_kernel<STAGES, KIND, MINB, NM><<<grid, 128, smem, st>>>(p);
    return cudaGetLastError();
}

// warp-strip sweep over groups of `members`/4 consecutive full regular strips whose members' S/N neighbours are consecutive elements:
// p.strip_reg holds the groups' StripReg records back to back, p.strip_count = number of groups; the map4_* tensor maps must
// have boxes of 32/members slabs x members elements
cudaError_t launch_sweep2_wstrip(int mode, const SweepParams &p, cudaStream_t st, int members)
{
    // shape A/B (BWK_WSTRIP_SHAPE): 4 = 2-stage ring, one row set, 128 registers: four CTAs (16 warps) per SM (4 members only);
    // 3 = 3-stage ring, two row sets, 168 registers: three CTAs per SM
    static const int shape = std::getenv("BWK_WSTRIP_SHAPE") ? std::atoi(std::getenv("BWK_WSTRIP_SHAPE")) : 3;
    if (shape == 4 && members == 4) {
        if (mode == MODE_SWEEP2) return launch_wstrip_t<2, SK_LAP, 4, 4>(p, st);
        if (mode == MODE_DSS) return launch_wstrip_t<2, SK_DSS, 4, 4>(p, st);
        if (mode == MODE_SWEEP2_POST) return launch_wstrip_t<2, SK_POST, 4, 4>(p, st);
    }
#define BWK_WS_CASE(NM_)                                                                   \
    if (members == NM_) {                                                                  \
        if (mode == MODE_SWEEP2) return launch_wstrip_t<3, SK_LAP, 3, NM_>(p, st);         \
        if (mode == MODE_DSS) return launch_wstrip_t<3, SK_DSS, 3, NM_>(p, st);            \
    }
    BWK_WS_CASE(4)
    BWK_WS_CASE(8)
    BWK_WS_CASE(16)
#undef BWK_WS_CASE
    if (mode == MODE_SWEEP2_POST && members == 4) return launch_wstrip_t<3, SK_POST, 3, 4>(p, st);
    if (mode == MODE_SWEEP2_POST && members == 8) return launch_wstrip_t<3, SK_POST, 3, 8>(p, st);
    return cudaErrorInvalidValue;
}

bool sweep1_pre_fits(int nlev) { return sweep_smem_bytes<3, 4>() + nlev * 128 <= 227 * 1024; }

cudaError_t launch_sweep(int mode, const SweepParams &p, cudaStream_t st)
{
    switch (mode) {
    case MODE_LAPLACE: return launch_sweep_t<MODE_LAPLACE, BWK_CFG_LAPLACE>(p, st);
    case MODE_SWEEP1: return launch_sweep_t<MODE_SWEEP1, BWK_CFG_SWEEP1>(p, st);
    case MODE_SWEEP1_PRE: return launch_sweep_t<MODE_SWEEP1_PRE, BWK_CFG_SWEEP1>(p, st);
    case MODE_SWEEP2: return launch_strip_t<BWK_CFG_SWEEP2, SK_LAP>(p, st);
    case MODE_DSS: return launch_strip_t<BWK_CFG_SWEEP2, SK_DSS>(p, st);
    case MODE_SWEEP2_POST: return launch_strip_t<BWK_CFG_SWEEP2, SK_POST>(p, st);
    default: return cudaErrorInvalidValue;
    }
}

// ------------------------------------------------------------------------------------------
// advance_hypervis_scalar, pointwise part (upstream HOMME prim_advection; not in the reference):
//     out = limiter2d_zero(Qdp * spheremp - dt*nu_q * Qtens) * rspheremp
// (the trailing rspheremp is the DSS scaling, folded into the producer as in sweep 1).  Fully coalesced:
// a thread owns one 16-byte chunk (2 GLL points), 8 consecutive lanes own one (np x np) level, the
// limiter's two sums per level are 3-step xor-shuffle reductions inside those 8 lanes.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) hypervis_post_kernel(const double *__restrict__ qdp, const double *__restrict__ qtens, double *__restrict__ out,
                                                            const double *__restrict__ elem, const double *__restrict__ rsph,
                                                            const int32_t *__restrict__ list, int first, int64_t nchunks, int kq, double coef, int limiter)
{
    const int64_t g = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const bool live = g < nchunks; // nchunks is a multiple of 8: the 8 lanes of a level are live together
    const int64_t slab = live ? g >> 3 : 0;
    const int c = (int)(g & 7);
    const int64_t li = slab / kq;
    const int le = list ? list[li] : first + (int)li;
    const size_t off = ((size_t)le * kq + (size_t)(slab - li * kq)) * 16 + 2 * c;
    const double2 a = *reinterpret_cast<const double2 *>(qdp + off), t = *reinterpret_cast<const double2 *>(qtens + off);
    const double2 mp = __ldg(reinterpret_cast<const double2 *>(elem + (size_t)le * ELEM_DOUBLES + 64 + 2 * c));
    const double2 rs = __ldg(reinterpret_cast<const double2 *>(rsph + (size_t)le * 16 + 2 * c));
    // same roundings as the CPU statement  Qdp*spheremp - (dt*nu_q)*Qtens  (no FMA contraction)
    double x = __dsub_rn(__dmul_rn(a.x, mp.x), __dmul_rn(coef, t.x));
    double y = __dsub_rn(__dmul_rn(a.y, mp.y), __dmul_rn(coef, t.y));
    if (limiter) { // limiter2d_zero on this level: clip negatives, rescale so that the level's sum is unchanged
        double mass = x + y;
        mass += __shfl_xor_sync(0xffffffffu, mass, 1);
        mass += __shfl_xor_sync(0xffffffffu, mass, 2);
        mass += __shfl_xor_sync(0xffffffffu, mass, 4);
        const bool neg = mass < 0.0;
        if (neg) { x = -x; y = -y; }
        x = x < 0.0 ? 0.0 : x;
        y = y < 0.0 ? 0.0 : y;
        double added = x + y;
        added += __shfl_xor_sync(0xffffffffu, added, 1);
        added += __shfl_xor_sync(0xffffffffu, added, 2);
        added += __shfl_xor_sync(0xffffffffu, added, 4);
        if (added > 0.0) {
            const double scale = fabs(mass) / added;
            x *= scale;
            y *= scale;
        }
        if (neg) { x = -x; y = -y; }
    }
    if (live) *reinterpret_cast<double2 *>(out + off) = make_double2(x * rs.x, y * rs.y);
}

cudaError_t launch_hypervis_post(const double *qdp, const double *qtens, double *out, const double *elem, const double *rsph,
                                 const int32_t *list, int first, int count, int kq, double coef, int limiter, cudaStream_t st)
{
    if (count <= 0) return cudaSuccess;
    const int64_t nchunks = (int64_t)count * kq * 8;
    hypervis_post_kernel<<<(unsigned)((nchunks + 255) / 256), 256, 0, st>>>(qdp, qtens, out, elem, rsph, list, first, nchunks, kq, coef, limiter);
    return cudaGetLastError();
}

__global__ void ratio_kernel(const double *__restrict__ num, const double *__restrict__ den, double *__restrict__ ratio, int64_t n)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) ratio[i] = num[i] / den[i];
}
// FP64 issue-rate microbenchmark (there is no FP64 peak in MEASURED_PEAKS.json, so the pipe utilisation
// ncu reports is put against a measured DFMA rate): 8 independent FMA chains per thread, 4096 FMAs each.
__global__ void __launch_bounds__(256) dfma_peak_kernel(double *out, double a, double b)
{
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i;
    for (int it = 0; it < 4096; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = fma(x[i], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i];
    if (s == 12345.678) out[0] = s; // never true: keeps the chains alive
}
cudaError_t launch_dfma_peak(double *out, int blocks, cudaStream_t st)
{
    dfma_peak_kernel<<<blocks, 256, 0, st>>>(out, 0.999999, 1e-9);
    return cudaGetLastError();
}

cudaError_t launch_ratio(const double *dpdiss, const double *dp, double *ratio, int64_t n, cudaStream_t st)
{
    ratio_kernel<<<148 * 8, 256, 0, st>>>(dpdiss, dp, ratio, n);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Plain variant of MODE_LAPLACE: one thread per slab, 128-bit global loads/stores, metric via
// the read-only path.  No staging: each warp-wide load touches 32 different 128-byte lines.
// Kept only as the ncu A/B baseline that justifies the TMA/swizzle staging above.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) laplace_plain_kernel(double *__restrict__ qtens, const double *__restrict__ metric, int elem_first, int kq)
{
    const int le = elem_first + blockIdx.y;
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= kq) return;
    double2 *slab = reinterpret_cast<double2 *>(qtens + ((size_t)le * kq + k) * 16);
    double v[16];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const double2 t = slab[c];
        v[2 * c] = t.x;
        v[2 * c + 1] = t.y;
    }
    laplace_slab(v, GlobalMetric{metric + (size_t)le * METRIC_DOUBLES});
#pragma unroll
    for (int c = 0; c < 8; ++c) slab[c] = make_double2(v[2 * c], v[2 * c + 1]);
}

cudaError_t launch_laplace_plain(double *qtens, const double *metric, int elem_first, int elem_count, int kq, cudaStream_t st)
{
    if (elem_count <= 0) return cudaSuccess;
    dim3 grid((unsigned)((kq + 127) / 128), (unsigned)elem_count);
    if (elem_count > 65535) {
        // grid.y limit: split the launch
        for (int e = 0; e < elem_count; e += 65535) {
            const int n = elem_count - e < 65535 ? elem_count - e : 65535;
            grid.y = (unsigned)n;
            laplace_plain_kernel<<<grid, 128, 0, st>>>(qtens, metric, elem_first + e, kq);
        }
        return cudaGetLastError();
    }
    laplace_plain_kernel<<<grid, 128, 0, st>>>(qtens, metric, elem_first, kq);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// biharmonic_wk_dp3d (upstream HOMME): the element-local part of one application, for the four
// fields of a level at once.  Field layout of an element = HOMME's edge-buffer order: T at slab k, v at nlev + m + 2k
// (m = lat-lon component), dp3d at 3 nlev + k.  One thread per (level, element):
//   T, dp3d : laplace_sphere_wk
//   v       : vlaplace_sphere_wk in the Cartesian form -- three Cartesian components through vec_sphere2cart(np,np,3,2),
//             laplace_sphere_wk on each, back with the transpose, + 2 spheremp v rrearth^2 (rigid rotation is not damped)
// PRODUCE: src -> dst, scaled by rspheremp (what the DSS strip sweep expects); otherwise in place and unscaled.
// Slabs move between global and shared memory warp-cooperatively (coalesced, cp.async on the way in); the Cartesian
// components live in shared memory around the three Laplacians, which keeps the kernel at ~128 registers.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}

// shared memory: element constants [192 doubles], then five 128-byte slots per level; 16-byte chunk c of slot r of the level
// owned by thread o sits at double2 index (r * 8 + c) * cs + o, cs = slots + 4: the owner's accesses are conflict-free, and
// so are the warp-cooperative copies below (chunk stride = 64 bytes mod 128).  A warp's 32 levels are contiguous in
// global memory in each of the three regions (T, v, dp3d), so the warp copies them in and out with fully coalesced 16-byte
// accesses (cp.async in, all 2 KB of a warp in flight before any arithmetic); slots 0..3 hold T, dp3d, v(1), v(2), the
// Cartesian components take slots 0, 1 (free once the scalars are stored) and 4.
template <bool PRODUCE>
__global__ void __launch_bounds__(128) dp3d_laplace_kernel(const double *src, double *dst, const double *__restrict__ metric,
                                                           const double *__restrict__ v2c, const double *__restrict__ elem,
                                                           const double *__restrict__ rsph, const int32_t *__restrict__ list, int elem_first, int nlev,
                                                           int nslot, int undamp)
{
    extern __shared__ __align__(16) double dp3d_s[];
    const int le = list ? list[blockIdx.x] : elem_first + (int)blockIdx.x;
    const int nt = blockDim.x, tid = threadIdx.x, lane = tid & 31, wbase = tid & ~31;
    const int kw = (int)blockIdx.y * nt + wbase;                 // first level of this warp
    const int nw = nlev - kw < 32 ? (nlev - kw > 0 ? nlev - kw : 0) : 32; // levels this warp owns (one per lane)
    const int cs = nslot + 4;
    double2 *slots = reinterpret_cast<double2 *>(dp3d_s + 192);
    const size_t ebase = (size_t)le * 4 * nlev * 16;
    const size_t off_t = ebase + (size_t)kw * 16, off_dp = ebase + ((size_t)3 * nlev + kw) * 16, off_v = ebase + ((size_t)nlev + 2 * (size_t)kw) * 16;
    if (nw > 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int g = i * 32 + lane;
            if (g < nw * 8) {
                cp_async16(smem_u32(slots + (0 * 8 + (g & 7)) * cs + wbase + (g >> 3)), src + off_t + 2 * g);
                cp_async16(smem_u32(slots + (1 * 8 + (g & 7)) * cs + wbase + (g >> 3)), src + off_dp + 2 * g);
            }
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int g = i * 32 + lane; // chunk of the warp's 2 nw contiguous v slabs: slab g/8 = component + 2 * level
            if (g < nw * 16) cp_async16(smem_u32(slots + ((2 + ((g >> 3) & 1)) * 8 + (g & 7)) * cs + wbase + (g >> 4)), src + off_v + 2 * g);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    // per-element constants through shared memory (uniform-address reads): folded metric, vec_sphere2cart, rspheremp, spheremp
    for (int t = tid; t < 192; t += nt)
        dp3d_s[t] = t < 64 ? metric[(size_t)le * METRIC_DOUBLES + t]
                  : t < 160 ? v2c[(size_t)le * 96 + (t - 64)]
                  : t < 176 ? rsph[(size_t)le * 16 + (t - 160)]
                            : elem[(size_t)le * ELEM_DOUBLES + 64 + (t - 176)];
    __syncthreads();
    if (nw == 0) return; // the whole warp
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    const bool active = lane < nw;
    const SmemMetric M{dp3d_s};
    const double *E = dp3d_s + 64, *rs = dp3d_s + 160, *sp = dp3d_s + 176;
    auto slot = [&](int r, int c) -> double2 & { return slots[(r * 8 + c) * cs + tid]; };
    if (active) {
#pragma unroll 1
        for (int f = 0; f < 2; ++f) {
            asm volatile("" ::: "memory"); // keeps the loop-invariant metric reads inside the loop (hoisted, they cost 128 registers)
            double s[16];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const double2 t = slot(f, c);
                s[2 * c] = t.x;
                s[2 * c + 1] = t.y;
            }
            laplace_slab(s, M);
            if (PRODUCE) {
#pragma unroll
                for (int t = 0; t < 16; ++t) s[t] *= rs[t];
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) slot(f, c) = make_double2(s[2 * c], s[2 * c + 1]);
        }
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int g = i * 32 + lane;
        if (g < nw * 8) {
            reinterpret_cast<double2 *>(dst + off_t)[g] = slots[(0 * 8 + (g & 7)) * cs + wbase + (g >> 3)];
            reinterpret_cast<double2 *>(dst + off_dp)[g] = slots[(1 * 8 + (g & 7)) * cs + wbase + (g >> 3)];
        }
    }
    __syncwarp(); // slots 0 and 1 are reused below
    if (active) {
        // vector: lat-lon -> three Cartesian components (slots 0, 1, 4), the Laplacian of each in place, then back
#pragma unroll 2
        for (int h = 0; h < 8; ++h) {
            const double2 u = slot(2, h), v = slot(3, h);
#pragma unroll
            for (int c = 0; c < 3; ++c)
                slot(c < 2 ? c : 4, h) = make_double2(fma(E[2 * h + 16 * c + 48], v.x, E[2 * h + 16 * c] * u.x),
                                                      fma(E[2 * h + 1 + 16 * c + 48], v.y, E[2 * h + 1 + 16 * c] * u.y));
        }
#pragma unroll 1
        for (int c = 0; c < 3; ++c) {
            asm volatile("" ::: "memory");
            const int r = c < 2 ? c : 4;
            double s[16];
#pragma unroll
            for (int h = 0; h < 8; ++h) {
                const double2 t = slot(r, h);
                s[2 * h] = t.x;
                s[2 * h + 1] = t.y;
            }
            laplace_slab(s, M);
#pragma unroll
            for (int h = 0; h < 8; ++h) slot(r, h) = make_double2(s[2 * h], s[2 * h + 1]);
        }
        const double und = undamp ? 2.0 * (BWK_RREARTH * BWK_RREARTH) : 0.0;
#pragma unroll 2
        for (int h = 0; h < 8; ++h) {
            const double2 uu = slot(2, h), vv = slot(3, h);
            const double2 l0 = slot(0, h), l1 = slot(1, h), l2 = slot(4, h);
            double o[4]; // u(2h), u(2h+1), v(2h), v(2h+1)
#pragma unroll
            for (int x = 0; x < 2; ++x) {
                const int t = 2 * h + x;
                const double w = und * sp[t];
                double ou = w * (x ? uu.y : uu.x), ov = w * (x ? vv.y : vv.x);
                const double l[3] = {x ? l0.y : l0.x, x ? l1.y : l1.x, x ? l2.y : l2.x};
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    ou = fma(l[c], E[t + 16 * c], ou);
                    ov = fma(l[c], E[t + 16 * c + 48], ov);
                }
                if (PRODUCE) {
                    ou *= rs[t];
                    ov *= rs[t];
                }
                o[x] = ou;
                o[2 + x] = ov;
            }
            slot(2, h) = make_double2(o[0], o[1]);
            slot(3, h) = make_double2(o[2], o[3]);
        }
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int g = i * 32 + lane;
        if (g < nw * 16) reinterpret_cast<double2 *>(dst + off_v)[g] = slots[((2 + ((g >> 3) & 1)) * 8 + (g & 7)) * cs + wbase + (g >> 4)];
    }
}

cudaError_t launch_dp3d_laplace(bool produce, const double *src, double *dst, const double *metric, const double *v2c, const double *elem,
                                const double *rsph, const int32_t *list, int elem_first, int elem_count, int nlev, int undamp, cudaStream_t st)
{
    if (elem_count <= 0) return cudaSuccess;
    const int threads = nlev >= 128 ? 128 : ((nlev + 31) / 32) * 32;
    const int nslot = nlev >= 128 ? 128 : ((nlev + 7) / 8) * 8; // slots only for the levels a block really has (72 -> 4 blocks per SM)
    const size_t smem = (192 + (size_t)5 * 8 * (nslot + 4) * 2) * sizeof(double);
    const dim3 grid((unsigned)elem_count, (unsigned)((nlev + threads - 1) / threads));
    static bool attr_set = false;
    if (!attr_set) {
        const int max_smem = (192 + 5 * 8 * (128 + 4) * 2) * (int)sizeof(double); // above the 48 KB default
        cudaError_t e = cudaFuncSetAttribute(dp3d_laplace_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(dp3d_laplace_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    if (produce)
        dp3d_laplace_kernel<true><<<grid, threads, smem, st>>>(src, dst, metric, v2c, elem, rsph, list, elem_first, nlev, nslot, undamp);
    else
        dp3d_laplace_kernel<false><<<grid, threads, smem, st>>>(src, dst, metric, v2c, elem, rsph, list, elem_first, nlev, nslot, undamp);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Setup / harness kernels
// ------------------------------------------------------------------------------------------
cudaError_t set_dvv(const double *dvv16_host, cudaStream_t st)
{
    return cudaMemcpyToSymbolAsync(c_dvv, dvv16_host, 16 * sizeof(double), 0, cudaMemcpyHostToDevice, st);
}

// One thread per (element, GLL point).  element_t layout F90:23-27 flattened.
__global__ void fold_metric_kernel(const double *__restrict__ elem, double *__restrict__ metric, int64_t nelem, int use_tensor)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nelem * 16) return;
    const int64_t e = g >> 4;
    const int pt = (int)(g & 15);
    fold_metric_point(elem + e * ELEM_DOUBLES, pt, use_tensor, metric + e * METRIC_DOUBLES + 4 * pt);
}

cudaError_t launch_fold_metric(const double *elem, double *metric, int64_t nelem, cudaStream_t st, int use_tensor)
{
    if (nelem <= 0) return cudaSuccess;
    const int64_t n = nelem * 16;
    fold_metric_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(elem, metric, nelem, use_tensor);
    return cudaGetLastError();
}

// HOMME edgeVpack analogue for the rows that leave this GPU.  Thread = one 16-byte half row.
__global__ void edge_pack_kernel(const double *__restrict__ tmp, const int32_t *__restrict__ row_elem, const int32_t *__restrict__ row_id, int kq,
                                 double *__restrict__ sendbuf)
{
    const int r = blockIdx.y;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= 2 * kq) return;
    const int slab = x >> 1, half = x & 1;
    const int64_t e = row_elem[r];
    const int row = row_id[r];
    const double *sl = tmp + (e * kq + slab) * 16;
    double2 v;
    if (row < 4) v = *reinterpret_cast<const double2 *>(sl + row * 4 + half * 2); // natural j-row
    else { // W (i=1) / E (i=np) column: stride np
        const int i = row == ROW_W ? 0 : 3;
        v = make_double2(sl[i + 4 * (2 * half)], sl[i + 4 * (2 * half + 1)]);
    }
    *reinterpret_cast<double2 *>(sendbuf + (((int64_t)r * kq + slab) * 4 + half * 2)) = v;
}

cudaError_t launch_edge_pack(const double *tmp, const int32_t *row_elem, const int32_t *row_id, int64_t nrows, int kq, double *sendbuf,
                             cudaStream_t st)
{
    for (int64_t r0 = 0; r0 < nrows; r0 += 65535) {
        const int64_t n = nrows - r0 < 65535 ? nrows - r0 : 65535;
        dim3 grid((unsigned)((2 * kq + 255) / 256), (unsigned)n);
        edge_pack_kernel<<<grid, 256, 0, st>>>(tmp, row_elem + r0, row_id + r0, kq, sendbuf + r0 * kq * 4);
    }
    return cudaGetLastError();
}

__device__ __forceinline__ double splitmix_unit(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    x ^= x >> 31;
    return (double)(x >> 11) * 0x1.0p-53;
}

// initialize_data (F90:48-58) by closed form: draw d (0-based) = table[d mod 131072].
__global__ void init_data_kernel(const double *__restrict__ table, int64_t nelemd_global, int64_t elem_first, int64_t nelem, int kq, int seed_mode,
                                 double *__restrict__ elem, double *__restrict__ qtens)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t per_elem = (int64_t)16 * kq;
    const int64_t n_elem_d = nelem * ELEM_DOUBLES, n_q = nelem * per_elem;
    const int64_t e_base = 16 + ELEM_DOUBLES * elem_first;
    const int64_t q_base = 16 + ELEM_DOUBLES * nelemd_global + per_elem * elem_first;
    for (int64_t t = tid; t < n_elem_d; t += stride)
        elem[t] = seed_mode == 0 ? table[(e_base + t) & 131071] : splitmix_unit((uint64_t)(e_base + t));
    for (int64_t t = tid; t < n_q; t += stride)
        qtens[t] = seed_mode == 0 ? table[(q_base + t) & 131071] : splitmix_unit((uint64_t)(q_base + t));
}

cudaError_t launch_init_data(const double *table, int64_t nelemd_global, int64_t elem_first, int64_t nelem, int kq, int seed_mode, double *elem,
                             double *qtens, cudaStream_t st)
{
    init_data_kernel<<<148 * 16, 256, 0, st>>>(table, nelemd_global, elem_first, nelem, kq, seed_mode, elem, qtens);
    return cudaGetLastError();
}

template <int NV>
__device__ __forceinline__ void block_reduce_store(double (&acc)[NV], double *out, bool is_max)
{
    __shared__ double sh[NV][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        double a = acc[v];
        for (int o = 16; o > 0; o >>= 1) {
            const double b = __shfl_down_sync(0xffffffffu, a, o);
            a = is_max ? fmax(a, b) : a + b;
        }
        if (lane == 0) sh[v][warp] = a;
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            double a = lane < (int)(blockDim.x >> 5) ? sh[v][lane] : 0.0;
            for (int o = 16; o > 0; o >>= 1) {
                const double b = __shfl_down_sync(0xffffffffu, a, o);
                a = is_max ? fmax(a, b) : a + b;
            }
            if (lane == 0) out[(size_t)v * gridDim.x + blockIdx.x] = a;
        }
    }
}

// partials[0..nb) = sum (q-qs)^2 ; partials[nb..2nb) = sum qs^2          (compute_l2norm, F90:69-73)
__global__ void l2_partials_kernel(const double *__restrict__ q, const double *__restrict__ qs, int64_t n, double *partials)
{
    double acc[2] = {0.0, 0.0};
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
        const double d = q[t] - qs[t];
        acc[0] += d * d;
        acc[1] += qs[t] * qs[t];
    }
    block_reduce_store<2>(acc, partials, false);
}
__global__ void sum_partials_kernel(const double *__restrict__ q, int64_t n, double *partials)
{
    double acc[2] = {0.0, 0.0};
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
        acc[0] += q[t];
        acc[1] += q[t] * q[t];
    }
    block_reduce_store<2>(acc, partials, false);
}
__global__ void period_defect_kernel(const double *__restrict__ q, int64_t n, int64_t period, double *partials)
{
    double acc[1] = {0.0};
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
        const double d = fabs(q[t] - q[t % period]);
        acc[0] = fmax(acc[0], d == d ? d : INFINITY);
    }
    block_reduce_store<1>(acc, partials, true);
}

// Partition-independent checksum of the BITS of the field: sum over values of mix64(bits) * (2*global_index + 1) mod 2^64.
// Integer addition is associative and commutative, so the per-rank sums of any contiguous partition add up to the 1-rank
// value exactly: equal checksums <=> (up to 2^-64) bit-identical fields in identical global positions.
__global__ void bits_checksum_kernel(const double *__restrict__ q, int64_t n, uint64_t index0, unsigned long long *partials)
{
    unsigned long long acc = 0;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
        uint64_t x = (uint64_t)__double_as_longlong(q[t]);
        x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
        x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
        x ^= x >> 31;
        acc += x * (2ull * (index0 + (uint64_t)t) + 1ull);
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    __shared__ unsigned long long sh[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) sh[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long a = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) a += sh[w];
        partials[blockIdx.x] = a;
    }
}
cudaError_t launch_bits_checksum(const double *q, int64_t n, uint64_t index0, unsigned long long *partials, int nblocks, cudaStream_t st)
{
    bits_checksum_kernel<<<nblocks, 256, 0, st>>>(q, n, index0, partials);
    return cudaGetLastError();
}

cudaError_t launch_l2_partials(const double *q, const double *qs, int64_t n, double *partials, int nblocks, cudaStream_t st)
{
    l2_partials_kernel<<<nblocks, 256, 0, st>>>(q, qs, n, partials);
    return cudaGetLastError();
}
cudaError_t launch_sum_partials(const double *q, int64_t n, double *partials, int nblocks, cudaStream_t st)
{
    sum_partials_kernel<<<nblocks, 256, 0, st>>>(q, n, partials);
    return cudaGetLastError();
}
cudaError_t launch_period_defect(const double *q, int64_t n, int64_t period, double *partials, int nblocks, cudaStream_t st)
{
    period_defect_kernel<<<nblocks, 256, 0, st>>>(q, n, period, partials);
    return cudaGetLastError();
}

} // namespace bwk
