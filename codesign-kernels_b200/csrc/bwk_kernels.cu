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

#include <cstdio>
#include <cstdlib>

namespace bwk {

// c_dvv[i + 4*l] == Dvv(i+1,l+1) of type derivative_t (F90:19-21)
__constant__ double c_dvv[16];

// F90:14: rrearth is a default-real literal -> float32-rounded value 0x1.65e9f8p-23
#define BWK_RREARTH 0x1.65e9f8p-23

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
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *map, int c0, int c1, int c2, uint32_t src)
{
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3}], [%4];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(src)
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

// ------------------------------------------------------------------------------------------
// The operator on one register-resident slab.  M points at the element's folded metric,
// 4 doubles per GLL point: {Ma, Mb, Mc, Md} with w1 = Ma dx + Mb dy, w2 = Mc dx + Md dy.
// s[i + 4*j] = s(i+1,j+1).  Overwrites s with laplace_sphere_wk(s) (F90:164-182).
// ------------------------------------------------------------------------------------------
template <typename MetricLoader>
__device__ __forceinline__ void laplace_slab(double (&s)[16], MetricLoader M)
{
    double w1[16], w2[16];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // dx(i,j) = sum_l Dvv(l,i) s(l,j) ; dy(i,j) = sum_l Dvv(l,j) s(i,l)      (F90:234-237)
            double dx = c_dvv[0 + 4 * i] * s[0 + 4 * j];
            double dy = c_dvv[0 + 4 * j] * s[i + 4 * 0];
#pragma unroll
            for (int l = 1; l < 4; ++l) {
                dx = fma(c_dvv[l + 4 * i], s[l + 4 * j], dx);
                dy = fma(c_dvv[l + 4 * j], s[i + 4 * l], dy);
            }
            const double2 mab = M(i + 4 * j, 0);
            const double2 mcd = M(i + 4 * j, 1);
            w1[i + 4 * j] = fma(mab.y, dy, mab.x * dx);
            w2[i + 4 * j] = fma(mcd.y, dy, mcd.x * dx);
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // out(i,j) = sum_l w1(l,j) Dvv(i,l) + w2(i,l) Dvv(j,l)                   (F90:155-156, sign/metric folded into M)
            double o = w1[0 + 4 * j] * c_dvv[i + 4 * 0];
            o = fma(w2[i + 4 * 0], c_dvv[j + 4 * 0], o);
#pragma unroll
            for (int l = 1; l < 4; ++l) {
                o = fma(w1[l + 4 * j], c_dvv[i + 4 * l], o);
                o = fma(w2[i + 4 * l], c_dvv[j + 4 * l], o);
            }
            s[i + 4 * j] = o;
        }
    }
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
// Sweep kernel (all three modes share the warp-private TMA pipeline)
//
//   MODE_LAPLACE : qtens tile -> laplace -> qtens tile (in place)
//   MODE_SWEEP1  : qtens tile -> laplace -> tmp tile, plus the W/E edge pack (the i=1 / i=np
//                  columns of every slab, stored kq-contiguous so that a neighbour reads them as
//                  fully coalesced 32-byte rows) -- the on-device part of HOMME's edgeVpack
//   MODE_SWEEP2  : tmp tile + the neighbours' coincident edge/corner values (edgeVunpack)
//                  -> * rspheremp -> laplace -> qtens tile
//
// Sweep 2 reads each neighbour's facing edge row as ONE 256-bit load per thread (LDG.E.256),
// straight from where sweep 1 left it: rows j=1/j=np of the neighbour's slab in tmp, the W/E
// pack, or the halo buffer filled by ncclRecv.  The loads are issued before the wait on the
// tile's mbarrier, so their latency overlaps the TMA wait; 12 resident warps per SM hide the rest.
// (A first version staged these rows with 4-D TMA boxes of 32-byte rows; it was bound by the TMA
// request rate -- see profiles/.)
// ------------------------------------------------------------------------------------------
template <int MODE>
struct StageBytes {
    static constexpr int value = MODE == MODE_SWEEP1 ? 4096 + 2048 : 4096;
};

template <int MODE, int STAGES, int WARPS>
constexpr int sweep_smem_bytes()
{
    // 1024 alignment slack + stages + mbarriers + metric + rsph + dss
    return 1024 + WARPS * STAGES * StageBytes<MODE>::value + ((WARPS * STAGES * 8 + 15) / 16) * 16 + METRIC_DOUBLES * 8 + 16 * 8 + 64 + 64;
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
__device__ __forceinline__ void bulk_store_1d(void *gdst, uint32_t src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(src), "r"(bytes) : "memory");
}

template <int MODE, int STAGES, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) sweep_kernel(const __grid_constant__ SweepParams p)
{
    constexpr int SB = StageBytes<MODE>::value;
    extern __shared__ unsigned char smem_raw[];
    // keep the pointer derived from smem_raw so that the compiler emits LDS/STS, not generic LD/ST
    unsigned char *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char *stages = smem;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + WARPS * STAGES * SB);
    double *metric_s = reinterpret_cast<double *>(smem + WARPS * STAGES * SB + ((WARPS * STAGES * 8 + 15) / 16) * 16);
    double *rsph_s = metric_s + METRIC_DOUBLES;
    uint32_t *dss_s = reinterpret_cast<uint32_t *>(rsph_s + 16);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int le = p.elem_list ? p.elem_list[blockIdx.x] : p.elem_first + (int)blockIdx.x;

    // ---- per-CTA staging of the element's folded metric (and DSS descriptor, rspheremp)
    for (int t = threadIdx.x; t < METRIC_DOUBLES; t += WARPS * 32) metric_s[t] = p.metric[(size_t)le * METRIC_DOUBLES + t];
    if (MODE == MODE_SWEEP2) {
        if (threadIdx.x < 16) rsph_s[threadIdx.x] = p.rsph[(size_t)le * 16 + threadIdx.x];
        if (threadIdx.x >= 32 && threadIdx.x < 48) dss_s[threadIdx.x - 32] = reinterpret_cast<const uint32_t *>(p.dss + le)[threadIdx.x - 32];
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

    // ---- sweep 2: the element's DSS descriptor decoded once into CTA-uniform values:
    //      per source a base address and the byte stride between consecutive (k,q) slabs
    const unsigned char *src_ptr[8]; // 0..3 edge rows S,N,W,E ; 4..7 diagonal corner values
    int src_stride[8];
    bool edge_rev[4];
    int corner_perm[4], corner_n[4];
    if (MODE == MODE_SWEEP2) {
        const uint32_t w_ekind = dss_s[8], w_erow = dss_s[9], w_erev = dss_s[10], w_dkind = dss_s[11], w_drow = dss_s[12], w_dpos = dss_s[13];
        const size_t kq = (size_t)p.kq;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int src = (int)dss_s[e];
            const uint32_t kind = (w_ekind >> (8 * e)) & 0xffu, row = (w_erow >> (8 * e)) & 0xffu;
            edge_rev[e] = ((w_erev >> (8 * e)) & 0xffu) != 0;
            if (src < 0) { src_ptr[e] = nullptr; src_stride[e] = 0; }
            else if (kind == SRC_TMP) { src_ptr[e] = reinterpret_cast<const unsigned char *>(p.tmp + ((size_t)src * kq * 16 + row * 4)); src_stride[e] = 128; }
            else if (kind == SRC_WEPACK) { src_ptr[e] = reinterpret_cast<const unsigned char *>(p.wepack + ((size_t)src * 2 + row) * kq * 4); src_stride[e] = 32; }
            else { src_ptr[e] = reinterpret_cast<const unsigned char *>(p.halo + (size_t)src * kq * 4); src_stride[e] = 32; }
            const int dsrc = (int)dss_s[4 + e];
            const uint32_t dkind = (w_dkind >> (8 * e)) & 0xffu, drow = (w_drow >> (8 * e)) & 0xffu, dpos = (w_dpos >> (8 * e)) & 0xffu;
            src_stride[4 + e] = 32;
            if (dsrc < 0) src_ptr[4 + e] = nullptr;
            else if (dkind == SRC_WEPACK) src_ptr[4 + e] = reinterpret_cast<const unsigned char *>(p.wepack + ((size_t)dsrc * 2 + drow) * kq * 4 + dpos);
            else src_ptr[4 + e] = reinterpret_cast<const unsigned char *>(p.halo + (size_t)dsrc * kq * 4 + dpos);
            corner_perm[e] = (int)((dss_s[15] >> (8 * e)) & 0xffu);
            corner_n[e] = (int)((dss_s[14] >> (8 * e)) & 0xffu);
        }
    }

    auto issue_load = [&](int n, int s) {
        const int kq0 = (t0 + warp + n * WARPS) * TILE_SLABS;
        const uint32_t dst = stage0 + s * SB, bar = bar0 + s * 8;
        mbar_expect_tx(bar, 4096u);
        tma_load_3d(dst, MODE == MODE_SWEEP2 ? &p.map_tmp : &p.map_q, 0, kq0, le, bar);
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

        // ---- sweep 2: fetch the neighbours' edge rows / corner values for my slab (registers)
        double nb[4][4], dg[4];
        if (MODE == MODE_SWEEP2) {
            const bool live = kq0 + lane < p.kq;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                nb[e][0] = nb[e][1] = nb[e][2] = nb[e][3] = 0.0;
                if (src_ptr[e] != nullptr && live) ldg256(src_ptr[e] + (size_t)(kq0 + lane) * src_stride[e], nb[e][0], nb[e][1], nb[e][2], nb[e][3]);
                dg[e] = 0.0;
                if (src_ptr[4 + e] != nullptr && live) dg[e] = ldg64(src_ptr[4 + e] + (size_t)(kq0 + lane) * 32);
            }
        }

        mbar_wait(bar0 + s * 8, (uint32_t)((n / STAGES) & 1));

        double v[16];
        unsigned char *row = tile + lane * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const double2 t = *reinterpret_cast<const double2 *>(row + ((c ^ sw) << 4));
            v[2 * c] = t.x;
            v[2 * c + 1] = t.y;
        }

        if (MODE == MODE_SWEEP2) {
            // ---- DSS: add the neighbours' coincident values (edgeVunpack), then * rspheremp
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (edge_rev[e]) { // neighbour's edge runs the other way
                    double t = nb[e][0]; nb[e][0] = nb[e][3]; nb[e][3] = t;
                    t = nb[e][1]; nb[e][1] = nb[e][2]; nb[e][2] = t;
                }
            }
            // mid-edge points: exactly two contributors, a+b is order independent
            if (src_ptr[EDGE_S] != nullptr) { v[1] += nb[EDGE_S][1]; v[2] += nb[EDGE_S][2]; }
            if (src_ptr[EDGE_N] != nullptr) { v[13] += nb[EDGE_N][1]; v[14] += nb[EDGE_N][2]; }
            if (src_ptr[EDGE_W] != nullptr) { v[4] += nb[EDGE_W][1]; v[8] += nb[EDGE_W][2]; }
            if (src_ptr[EDGE_E] != nullptr) { v[7] += nb[EDGE_E][1]; v[11] += nb[EDGE_E][2]; }
            // corners: up to four contributors, summed in ascending global point order
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int pt = c == 0 ? 0 : (c == 1 ? 3 : (c == 2 ? 12 : 15));
                const double cv0 = v[pt];
                const double cv1 = nb[c < 2 ? EDGE_S : EDGE_N][(c & 1) ? 3 : 0];
                const double cv2 = nb[(c & 1) ? EDGE_E : EDGE_W][c < 2 ? 0 : 3];
                const double cv3 = dg[c];
                const int perm = corner_perm[c], cn = corner_n[c];
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
#pragma unroll
            for (int t = 0; t < 16; t += 2) {
                const double2 r = *reinterpret_cast<const double2 *>(rsph_s + t);
                v[t] *= r.x;
                v[t + 1] *= r.y;
            }
        }

        laplace_slab(v, M);

#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<double2 *>(row + ((c ^ sw) << 4)) = make_double2(v[2 * c], v[2 * c + 1]);
        if (MODE == MODE_SWEEP1) {
            // W/E edge pack staging: [side][lane][4]
            unsigned char *we = tile + 4096 + lane * 32;
            *reinterpret_cast<double2 *>(we) = make_double2(v[0], v[4]);
            *reinterpret_cast<double2 *>(we + 16) = make_double2(v[8], v[12]);
            *reinterpret_cast<double2 *>(we + 1024) = make_double2(v[3], v[7]);
            *reinterpret_cast<double2 *>(we + 1024 + 16) = make_double2(v[11], v[15]);
        }
        fence_proxy_async();
        __syncwarp();

        if (lane == 0) {
            const uint32_t src = stage0 + s * SB;
            tma_store_3d(MODE == MODE_SWEEP1 ? &p.map_tmp : &p.map_q, 0, kq0, le, src);
            if (MODE == MODE_SWEEP1) {
                const uint32_t nvalid = (uint32_t)min(TILE_SLABS, p.kq - kq0);
                double *w = p.wepack + (((size_t)le * 2) * p.kq + kq0) * 4;
                bulk_store_1d(w, src + 4096, nvalid * 32u);
                bulk_store_1d(w + (size_t)p.kq * 4, src + 4096 + 1024, nvalid * 32u);
            }
            bulk_commit();
            // refill the stage tile n-1 used, once its store has finished reading shared memory
            if (n >= 1 && n - 1 + STAGES < my_n) {
                bulk_wait_read<1>();
                issue_load(n - 1 + STAGES, (n - 1) % STAGES);
            }
        }
    }
    if (lane == 0) bulk_wait<0>();
}

// kernel configurations: (STAGES, WARPS, MINB)
#define BWK_CFG_LAPLACE 3, 4, 3
#define BWK_CFG_SWEEP1 3, 4, 3
#define BWK_CFG_SWEEP2 3, 4, 3

template <int MODE>
struct ModeWarps {
    static constexpr int value = 4;
};

SweepGeometry choose_geometry(int kq, int64_t nelem, int mode)
{
    (void)mode;
    const int warps = 4;
    const int ntiles = (kq + TILE_SLABS - 1) / TILE_SLABS;
    // enough CTAs for ~8 per SM, but at most ~6 tiles per warp (short CTAs balance better, and in
    // sweep 2 a short unit keeps a neighbour's rows L2-resident until its own CTA reads them)
    const int64_t target_ctas = 148 * 8;
    int64_t tpu = ((int64_t)ntiles * nelem + target_ctas - 1) / target_ctas;
    if (tpu < warps) tpu = warps;
    if (tpu > 6 * warps) tpu = 6 * warps;
    if (tpu > ntiles) tpu = ntiles;
    SweepGeometry g;
    g.units = (int)((ntiles + tpu - 1) / tpu);
    g.tiles_per_unit = (int)((ntiles + g.units - 1) / g.units); // rebalance
    return g;
}

template <int MODE, int STAGES, int WARPS, int MINB>
static cudaError_t launch_sweep_t(const SweepParams &p, cudaStream_t st)
{
    constexpr int smem = sweep_smem_bytes<MODE, STAGES, WARPS>();
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(sweep_kernel<MODE, STAGES, WARPS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (p.elem_count <= 0 || p.units <= 0) return cudaSuccess;
    dim3 grid((unsigned)p.elem_count, (unsigned)p.units);
    sweep_kernel<MODE, STAGES, WARPS, MINB><<<grid, WARPS * 32, smem, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_sweep(int mode, const SweepParams &p, cudaStream_t st)
{
    switch (mode) {
    case MODE_LAPLACE: return launch_sweep_t<MODE_LAPLACE, BWK_CFG_LAPLACE>(p, st);
    case MODE_SWEEP1: return launch_sweep_t<MODE_SWEEP1, BWK_CFG_SWEEP1>(p, st);
    case MODE_SWEEP2: return launch_sweep_t<MODE_SWEEP2, BWK_CFG_SWEEP2>(p, st);
    default: return cudaErrorInvalidValue;
    }
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
// Setup / harness kernels
// ------------------------------------------------------------------------------------------
cudaError_t set_dvv(const double *dvv16_host, cudaStream_t st)
{
    return cudaMemcpyToSymbolAsync(c_dvv, dvv16_host, 16 * sizeof(double), 0, cudaMemcpyHostToDevice, st);
}

// One thread per (element, GLL point).  element_t layout F90:23-27 flattened.
__global__ void fold_metric_kernel(const double *__restrict__ elem, double *__restrict__ metric, int64_t nelem)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nelem * 16) return;
    const int64_t e = g >> 4;
    const int pt = (int)(g & 15);
    const double *E = elem + e * ELEM_DOUBLES;
    const double D11 = E[pt], D21 = E[16 + pt], D12 = E[32 + pt], D22 = E[48 + pt]; // Dinv(i,j,c,a) at 16c+32a
    const double sp = E[64 + pt];
    const double T11 = E[80 + pt], T21 = E[96 + pt], T12 = E[112 + pt], T22 = E[128 + pt];
    // g = T * D^T * (dx,dy)  (F90:128-133 then F90:175-178)
    const double G1x = T11 * D11 + T12 * D12, G1y = T11 * D21 + T12 * D22;
    const double G2x = T21 * D11 + T22 * D12, G2y = T21 * D21 + T22 * D22;
    // vt = D * g  (F90:147-148), then * spheremp * rrearth (F90:155-156) and the gradient's rrearth (F90:124-125)
    const double f = -(BWK_RREARTH * BWK_RREARTH) * sp;
    double *M = metric + e * METRIC_DOUBLES + 4 * pt;
    M[0] = f * (D11 * G1x + D12 * G2x);
    M[1] = f * (D11 * G1y + D12 * G2y);
    M[2] = f * (D21 * G1x + D22 * G2x);
    M[3] = f * (D21 * G1y + D22 * G2y);
}

cudaError_t launch_fold_metric(const double *elem, double *metric, int64_t nelem, cudaStream_t st)
{
    if (nelem <= 0) return cudaSuccess;
    const int64_t n = nelem * 16;
    fold_metric_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(elem, metric, nelem);
    return cudaGetLastError();
}

// HOMME edgeVpack analogue for the rows that leave this GPU.  Thread = one 16-byte half row.
__global__ void edge_pack_kernel(const double *__restrict__ tmp, const double *__restrict__ wepack, const int32_t *__restrict__ row_elem,
                                 const int32_t *__restrict__ row_id, int kq, double *__restrict__ sendbuf)
{
    const int r = blockIdx.y;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= 2 * kq) return;
    const int slab = x >> 1, half = x & 1;
    const int64_t e = row_elem[r];
    const int row = row_id[r];
    const double *src = row < 4 ? tmp + ((e * kq + slab) * 16 + row * 4 + half * 2) : wepack + (((e * 2 + (row - 4)) * kq + slab) * 4 + half * 2);
    *reinterpret_cast<double2 *>(sendbuf + (((int64_t)r * kq + slab) * 4 + half * 2)) = *reinterpret_cast<const double2 *>(src);
}

cudaError_t launch_edge_pack(const double *tmp, const double *wepack, const int32_t *row_elem, const int32_t *row_id, int64_t nrows, int kq,
                             double *sendbuf, cudaStream_t st)
{
    for (int64_t r0 = 0; r0 < nrows; r0 += 65535) {
        const int64_t n = nrows - r0 < 65535 ? nrows - r0 : 65535;
        dim3 grid((unsigned)((2 * kq + 255) / 256), (unsigned)n);
        edge_pack_kernel<<<grid, 256, 0, st>>>(tmp, wepack, row_elem + r0, row_id + r0, kq, sendbuf + r0 * kq * 4);
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
