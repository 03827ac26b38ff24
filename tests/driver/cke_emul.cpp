// TEST INFRASTRUCTURE ONLY -- compiles the device source of libcke_b200.so
// (codesign-kernels_b200/csrc/cke_kernels.cuh, unmodified) for the HOST: the CUDA vocabulary is
// mapped onto blockDim.x host threads and a barrier, CTAs run one after another.  This is how the
// kernel's indexing, staging and operation order are checked against the oracle on a box
// without a GPU; it is not a CPU path of the product (nothing in the package builds or loads it).
#include <algorithm>
#include <barrier>
#include <cstddef>
#include <cstring>
#include <thread>
#include <vector>

struct emu_dim3 {
    unsigned x = 1, y = 1, z = 1;
};
static thread_local emu_dim3 threadIdx, blockIdx;
static emu_dim3 blockDim, gridDim;
static std::barrier<> *emu_barrier = nullptr;
static inline void __syncthreads() { emu_barrier->arrive_and_wait(); }

struct double2 {
    double x, y;
};
static inline double __dmul_rn(double a, double b) { return a * b; }  // built with -ffp-contract=off
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
template <typename T> static inline T __ldg(const T *p) { return *p; }
using std::min;

#define __global__
#define __device__
#define __forceinline__ inline
#define __launch_bounds__(n)
#define __align__(n) __attribute__((aligned(n)))
#define __shared__

namespace cke_dev {
alignas(16) unsigned char smem_raw[48 * 1024];
}
#include "../../codesign-kernels_b200/csrc/cke_kernels.cuh"

using namespace cke_dev;

template <typename F> static void launch(unsigned grid, unsigned block, F kernel) {
    blockDim.x = block;
    gridDim.x = grid;
    std::barrier<> bar((std::ptrdiff_t)block);
    emu_barrier = &bar;
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < block; ++t)
        pool.emplace_back([=, &bar] {
            threadIdx.x = t;
            for (unsigned b = 0; b < grid; ++b) {
                blockIdx.x = b;
                kernel();
                bar.arrive_and_wait();  // CTAs share the one shared-memory buffer: run them in turn
            }
        });
    for (auto &th : pool) th.join();
    emu_barrier = nullptr;
}

// One iteration of implementation `impl` exactly as cke_b200.cu's launch_iteration issues it,
// on host pointers.  tracerCur is masked in place by impl 1.  Returns 0, or 1 on bad arguments.
extern "C" int cke_emul_iteration(int unroll, int impl, int grid, int block, int edges_per_cta, int nEdges, int nCells,
                                  int nVertLevels, int nvldim, int nAdv, const int *nAdvCellsForEdge,
                                  const int *minLevelCell, const int *maxLevelCell, const int *advCellsForEdge,
                                  double *tracerCur, const double *normalThicknessFlux, const double *advMaskHighOrder,
                                  const double *cellMask, const double *advCoefs, const double *advCoefs3rd,
                                  double coef3rdOrder, double *highOrderFlx) {
    if ((unroll != 1 && unroll != 5 && unroll != 10) || (impl < 1 || impl > 3) || grid < 1 || block < edges_per_cta || edges_per_cta < 1) return 1;
    if ((size_t)edges_per_cta * nAdv * sizeof(Rec) + edges_per_cta * sizeof(int) > sizeof smem_raw) return 1;
    Problem p;
    p.nIters = 1; p.nEdges = nEdges; p.nCells = nCells; p.nVertLevels = nVertLevels; p.nvldim = nvldim; p.nAdv = nAdv;
    p.coef3rdOrder = coef3rdOrder;
    p.nAdvCellsForEdge = const_cast<int *>(nAdvCellsForEdge);
    p.minLevelCell = const_cast<int *>(minLevelCell);
    p.maxLevelCell = const_cast<int *>(maxLevelCell);
    p.advCellsForEdge = const_cast<int *>(advCellsForEdge);
    p.tracerCur = tracerCur;
    p.cellMask = const_cast<double *>(cellMask);
    p.normalThicknessFlux = const_cast<double *>(normalThicknessFlux);
    p.advMaskHighOrder = const_cast<double *>(advMaskHighOrder);
    p.advCoefs = const_cast<double *>(advCoefs);
    p.advCoefs3rd = const_cast<double *>(advCoefs3rd);
    p.highOrderFlx = highOrderFlx;
    p.range_no_mask = impl == 3 ? 1 : 0;  // impl 3: the original loop form (yakl_original_loop): level ranges, no cellMask factor
    const bool vec2 = nvldim % 2 == 0;
    if (impl == 1) {
        const size_t n = (size_t)nCells * nvldim;
        launch((unsigned)grid, (unsigned)block, [=] { mask_tracer_kernel(p.tracerCur, p.cellMask, n); });
    }
#define EMUL_CASE(V_, R_, U_)                                                                              \
    if (vec2 == (V_ == 2) && (impl >= 2) == R_ && unroll == U_)                                            \
        launch((unsigned)grid, (unsigned)block, [=] { flux_kernel<V_, R_, U_>(p, edges_per_cta); });
    EMUL_CASE(1, false, 1) EMUL_CASE(1, false, 5) EMUL_CASE(1, false, 10)
    EMUL_CASE(2, false, 1) EMUL_CASE(2, false, 5) EMUL_CASE(2, false, 10)
    EMUL_CASE(1, true, 1) EMUL_CASE(1, true, 5) EMUL_CASE(1, true, 10)
    EMUL_CASE(2, true, 1) EMUL_CASE(2, true, 5) EMUL_CASE(2, true, 10)
#undef EMUL_CASE
    return 0;
}
