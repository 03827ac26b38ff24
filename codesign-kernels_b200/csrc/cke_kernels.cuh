// cke_kernels.cuh -- device code of libcke_b200.so (see cke_b200.cu for the design notes).
// Kept in a header so that tests/driver/cke_emul.cpp can compile the SAME kernel source for the
// host (CUDA vocabulary mapped onto host threads + a barrier) and check it against the oracle on
// boxes without a GPU.  The product only ever includes it from cke_b200.cu under nvcc.
#ifndef CKE_KERNELS_CUH
#define CKE_KERNELS_CUH
#include <cstddef>

namespace cke_dev {

struct Problem {
    int nIters, nEdges, nCells, nVertLevels, nvldim, nAdv;
    double coef3rdOrder;
    int *nAdvCellsForEdge, *minLevelCell, *maxLevelCell, *advCellsForEdge;
    double *tracerCur, *cellMask, *normalThicknessFlux, *advMaskHighOrder, *advCoefs, *advCoefs3rd;
    double *highOrderFlx;
    int range_no_mask = 0;  // RANGE form without the cellMask factor: the reference's yakl_original_loop (nested_yakl_c.cpp:184-243)
};

constexpr int kThreads = 256;
constexpr int kMaxEdgesPerCta = 16;
constexpr int kCtasPerSm = 8;  // 256 threads x 8 = 2048 resident threads per SM

// per (edge, contributing cell) record staged in shared memory
struct __align__(16) Rec {
    double cp, cm;  // coef1 + coef3, coef1 - coef3
    int cell;       // 0-based row of tracerCur
    int kmin, kmax; // 0-based inclusive level range of that cell (RANGE form only)
    int pad;
};

template <int VEC> struct Vec;
template <> struct Vec<1> {
    using T = double;
    static __device__ __forceinline__ T ld(const double *p) { return *p; }
    static __device__ __forceinline__ T ldg(const double *p) { return __ldg(p); }
    static __device__ __forceinline__ void st(double *p, T v) { *p = v; }
    static __device__ __forceinline__ double get(const T &v, int) { return v; }
    static __device__ __forceinline__ void set(T &v, int, double x) { v = x; }
};
template <> struct Vec<2> {
    using T = double2;
    static __device__ __forceinline__ T ld(const double *p) { return *reinterpret_cast<const double2 *>(p); }
    static __device__ __forceinline__ T ldg(const double *p) { return __ldg(reinterpret_cast<const double2 *>(p)); }
    static __device__ __forceinline__ void st(double *p, T v) { *reinterpret_cast<double2 *>(p) = v; }
    static __device__ __forceinline__ double get(const T &v, int s) { return s ? v.y : v.x; }
    static __device__ __forceinline__ void set(T &v, int s, double x) {
        if (s) v.y = x; else v.x = x;
    }
};

// pass 1 of cke_impl1.cpp:12-17 (nested.F90:307-311): tracerCur *= cellMask, in place
__global__ void __launch_bounds__(kThreads) mask_tracer_kernel(double *__restrict__ tracer,
                                                               const double *__restrict__ mask, size_t n) {
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < n; t += (size_t)gridDim.x * blockDim.x)
        tracer[t] = __dmul_rn(tracer[t], mask[t]);
}

// RANGE = false: pass 2 of cke_impl1.cpp:18-36 (collapsed (edge, level) loop, nested.F90:332-349)
// RANGE = true : cke_impl2.cpp:27-68 at pack size 1 (per-cell level range, tracer*cellMask); with p.range_no_mask the
//                original loop form nested.F90:119-157 / nested_yakl_c.cpp:184-243 (per-cell level range, no mask factor)
template <int VEC, bool RANGE, int U>
__global__ void __launch_bounds__(kThreads) flux_kernel(const Problem p, const int edges_per_cta) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Rec *rec = reinterpret_cast<Rec *>(smem_raw);                  // [edges_per_cta][nAdv]
    int *nadv = reinterpret_cast<int *>(rec + (size_t)edges_per_cta * p.nAdv);  // [edges_per_cta]
    using V = Vec<VEC>;
    const unsigned groups = (unsigned)(p.nvldim / VEC);  // level groups per edge
    const int nblocks = (p.nEdges + edges_per_cta - 1) / edges_per_cta;

    for (int blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const int e0 = blk * edges_per_cta;
        const int ne = min(edges_per_cta, p.nEdges - e0);
        // ---- phase A: stage the edge records -------------------------------------------------
        for (int t = threadIdx.x; t < ne * p.nAdv; t += blockDim.x) {
            const int el = t / p.nAdv, i = t - el * p.nAdv;
            const size_t src = (size_t)(e0 + el) * p.nAdv + i;
            Rec r;
            r.cell = 0; r.kmin = 0; r.kmax = -1; r.cp = 0.0; r.cm = 0.0; r.pad = 0;
            if (i < p.nAdvCellsForEdge[e0 + el]) {
                const int c = p.advCellsForEdge[src] - 1;
                const double coef1 = p.advCoefs[src];
                const double coef3 = __dmul_rn(p.advCoefs3rd[src], p.coef3rdOrder);
                r.cell = c;
                r.cp = __dadd_rn(coef1, coef3);
                r.cm = __dsub_rn(coef1, coef3);
                if (RANGE) {
                    r.kmin = p.minLevelCell[c] - 1;
                    r.kmax = p.maxLevelCell[c] - 1;
                }
            }
            rec[t] = r;
        }
        if ((int)threadIdx.x < ne) nadv[threadIdx.x] = min(p.nAdvCellsForEdge[e0 + threadIdx.x], p.nAdv);
        __syncthreads();
        // ---- phase B: every (edge, level group) point of the block ---------------------------
        for (unsigned pt = threadIdx.x; pt < (unsigned)ne * groups; pt += blockDim.x) {
            const unsigned el = pt / groups, kg = pt - el * groups;
            const size_t row = (size_t)(e0 + el) * p.nvldim + (size_t)kg * VEC;
            const typename V::T ntf = V::ld(p.normalThicknessFlux + row);
            const typename V::T msk = V::ld(p.advMaskHighOrder + row);
            double coef2[VEC], acc[VEC];
            bool neg[VEC];
#pragma unroll
            for (int s = 0; s < VEC; ++s) {
                coef2[s] = __dmul_rn(V::get(ntf, s), V::get(msk, s));
                neg[s] = V::get(ntf, s) < 0.0;  // csgn = ntf < 0 ? -1 : 1, cke_impl1.cpp:23-26
                acc[s] = 0.0;
            }
            const Rec *er = rec + (size_t)el * p.nAdv;
            const int n = nadv[el];
            // U gathers in flight per thread: the records and tracer rows of a batch are all loaded
            // before the first is consumed (the sum itself stays in the reference order i = 1..n)
            for (int i0 = 0; i0 < n; i0 += U) {
                typename V::T tr[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const Rec *ru = er + min(i0 + u, n - 1);  // a clamped duplicate is loaded and ignored
                    const size_t trow = (size_t)ru->cell * p.nvldim + (size_t)kg * VEC;
                    if (RANGE) {
                        // rows of cells whose level range misses this group are not fetched at all
                        const int k0 = (int)kg * VEC;
                        if (k0 + VEC - 1 >= ru->kmin && k0 <= ru->kmax) {
                            tr[u] = V::ldg(p.tracerCur + trow);
                            if (!p.range_no_mask) {
                                const typename V::T cm = V::ldg(p.cellMask + trow);
#pragma unroll
                                for (int s = 0; s < VEC; ++s) V::set(tr[u], s, __dmul_rn(V::get(tr[u], s), V::get(cm, s)));
                            }
                        } else {
#pragma unroll
                            for (int s = 0; s < VEC; ++s) V::set(tr[u], s, 0.0);
                        }
                    } else {
                        tr[u] = V::ldg(p.tracerCur + trow);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (i0 + u < n) {
                        const Rec r = er[i0 + u];  // re-read from shared memory: cheaper than U records in registers
#pragma unroll
                        for (int s = 0; s < VEC; ++s) {
                            const double sel = neg[s] ? r.cm : r.cp;
                            const double term = __dmul_rn(__dmul_rn(V::get(tr[u], s), coef2[s]), sel);
                            if (RANGE) {
                                const int k = (int)kg * VEC + s;
                                if (k >= r.kmin && k <= r.kmax) acc[s] = __dadd_rn(acc[s], term);
                            } else {
                                acc[s] = __dadd_rn(acc[s], term);
                            }
                        }
                    }
                }
            }
            typename V::T out;
#pragma unroll
            for (int s = 0; s < VEC; ++s) V::set(out, s, acc[s]);
            V::st(p.highOrderFlx + row, out);
        }
        __syncthreads();  // the records are overwritten by the next block of edges
    }
}

}  // namespace cke_dev
#endif
