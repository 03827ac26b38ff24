// TEST INFRASTRUCTURE ONLY -- compiles the device source of libmpdata_b200.so
// (codesign-kernels_b200/csrc/mpdata_kernels.cuh, unmodified) for the HOST: the CUDA vocabulary is
// mapped onto blockDim.x host threads and a barrier, CTAs run one after another.  This is how the
// kernel's indexing, phase ordering (barrier-phased in-place updates in shared memory) and
// operation order are checked against the oracle on a box without a GPU; it is not a CPU path of
// the product (nothing in the package builds or loads it).
#include <barrier>
#include <cmath>
#include <cstddef>
#include <memory>
#include <thread>
#include <vector>

struct emu_dim3 {
    unsigned x = 1, y = 1, z = 1;
};
static thread_local emu_dim3 threadIdx, blockIdx;
static emu_dim3 blockDim, gridDim;
static std::barrier<> *emu_barrier = nullptr;
static inline void __syncthreads() { emu_barrier->arrive_and_wait(); }
static std::barrier<> *emu_warp_barrier[32] = {};  // one per warp of the CTA: __syncwarp() of the marching form
static inline void __syncwarp() { emu_warp_barrier[threadIdx.x >> 5]->arrive_and_wait(); }

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(a, b)
#define __align__(n) __attribute__((aligned(n)))
#define __shared__

namespace mpd_dev {
alignas(16) unsigned char smem_raw[227 * 1024];  // the block-scope `extern ... smem_raw[]` of the kernel binds to this
}
#include "../../codesign-kernels_b200/csrc/mpdata_kernels.cuh"
#ifndef MPDATA_EMUL_NO_MARCH  // (the race check's negative control swaps mpdata_kernels.cuh for a modified copy and leaves this form out)
#include "../../codesign-kernels_b200/csrc/mpdata_march.cuh"
#endif

// the kernels' quotient of the limiter ratios (mpdata_kernels.cuh: div_ratio), for the test against the IEEE quotient
extern "C" double mpdata_emul_div_ratio(double num, double den) { return mpd_dev::div_ratio(num, den); }

// 1: use the run-time-size kernel even at the reference sizes (the library's MPDATA_B200_GENERIC=1)
static int mpdata_emul_generic = 0;
extern "C" void mpdata_emul_set_generic(int on) { mpdata_emul_generic = on; }

// One step exactly as mpdata_b200.cu's launch_step issues it (grid CTAs of `block` threads, `chunks`
// level ranges per slice), on
// host pointers.  Returns 0, 1 on bad arguments, 3 if a slice does not fit the shared memory.
// 1: the prefetching form (second set of input planes, the next item's loads issued one item ahead)
static int mpdata_emul_prefetch = 0;
extern "C" void mpdata_emul_set_prefetch(int on) { mpdata_emul_prefetch = on; }

extern "C" int mpdata_emul_step(int grid, int block, int chunks, int ns, int nx, int nz, double *f, const double *u, const double *w,
                                const double *rho, const double *rhow, const double *adz, double *flux) {
    if (grid < 1 || block < 1 || block > 1024 || ns < 1 || nx < 1 || nz < 3 || chunks < 1 || chunks > nz - 1) return 1;
    const int nlmax = mpd_dev::max_local_levels(nz - 1, chunks);
    if ((mpd_dev::item_doubles(nx, nlmax) + (mpdata_emul_prefetch ? mpd_dev::prefetch_doubles(nx, nlmax) : 0)) * sizeof(double) > sizeof mpd_dev::smem_raw) return 3;
    const int prefetch = mpdata_emul_prefetch;
    for (size_t j = 0; j < sizeof mpd_dev::smem_raw; ++j) mpd_dev::smem_raw[j] = 0xff;  // shared memory starts uninitialised: all-ones bytes = NaN, which would poison any output it reached
    const std::vector<double> fin(f, f + (size_t)ns * (nx + 6) * (nz - 1));  // the library ping-pongs two device buffers
    mpd_dev::Problem p{ns, nx, nz, fin.data(), f, u, w, rho, rhow, adz, flux};
    blockDim.x = (unsigned)block;
    gridDim.x = (unsigned)grid;
    std::barrier<> bar((std::ptrdiff_t)block);
    emu_barrier = &bar;
    std::vector<std::thread> pool;
    for (int t = 0; t < block; ++t)
        pool.emplace_back([=, &bar] {
            threadIdx.x = (unsigned)t;
            for (int b = 0; b < grid; ++b) {
                blockIdx.x = (unsigned)b;
                // the library's dispatch: the reference sizes run the instantiation with compile-time plane sizes
                if (nx == 32 && nz == 58 && !mpdata_emul_generic) mpd_dev::advect_kernel<32, 58>(p, chunks, nlmax, prefetch);
                else mpd_dev::advect_kernel<0, 0>(p, chunks, nlmax, prefetch);
                bar.arrive_and_wait();  // CTAs share the one shared-memory buffer: run them in turn
            }
        });
    for (auto &th : pool) th.join();
    emu_barrier = nullptr;
    return 0;
}

#ifndef MPDATA_EMUL_NO_MARCH
// One step of the marching form (mpdata_march.cuh: a warp per slice, no CTA-wide barrier) as mpdata_b200.cu launches it:
// grid CTAs of `block` threads (a multiple of 32).  Compiled for the reference sizes and for a few small planes the tests
// use (sizes are template parameters of that kernel).  Returns 0, 1 on bad arguments, 2 if the sizes are not instantiated,
// 3 if the CTA's rings do not fit the shared memory.
template <int NX, int NZ, int NS2>
static void march_cta(const mpd_dev::Problem &p) { mpd_dev::advect_march_kernel<NX, NZ, NS2>(p); }
// slices_per_warp: 1 or 2 (the kernel's NS2)
extern "C" int mpdata_emul_march_step(int grid, int block, int slices_per_warp, int ns, int nx, int nz, double *f, const double *u, const double *w,
                                      const double *rho, const double *rhow, const double *adz, double *flux) {
    if (grid < 1 || block < 32 || block > 1024 || block % 32 || ns < 1 || slices_per_warp < 1 || slices_per_warp > 2) return 1;
    void (*cta)(const mpd_dev::Problem &) = nullptr;
    const bool two = slices_per_warp == 2;
    if (nx == 32 && nz == 58) cta = two ? march_cta<32, 58, 2> : march_cta<32, 58, 1>;
    else if (nx == 5 && nz == 8) cta = two ? march_cta<5, 8, 2> : march_cta<5, 8, 1>;
    else if (nx == 7 && nz == 12) cta = two ? march_cta<7, 12, 2> : march_cta<7, 12, 1>;
    else if (nx == 1 && nz == 3) cta = two ? march_cta<1, 3, 2> : march_cta<1, 3, 1>;
    else if (nx == 40 && nz == 6) cta = two ? march_cta<40, 6, 2> : march_cta<40, 6, 1>;
    if (!cta) return 2;
    if ((size_t)(block / 32) * slices_per_warp * mpd_dev::march_warp_doubles(nx) * sizeof(double) > sizeof mpd_dev::smem_raw) return 3;
    for (size_t j = 0; j < sizeof mpd_dev::smem_raw; ++j) mpd_dev::smem_raw[j] = 0xff;  // NaN: an uninitialised read would poison the output
    const std::vector<double> fin(f, f + (size_t)ns * (nx + 6) * (nz - 1));
    mpd_dev::Problem p{ns, nx, nz, fin.data(), f, u, w, rho, rhow, adz, flux};
    blockDim.x = (unsigned)block;
    gridDim.x = (unsigned)grid;
    std::barrier<> bar((std::ptrdiff_t)block);
    std::vector<std::unique_ptr<std::barrier<>>> wb;
    for (int wi = 0; wi < block / 32; ++wi) {
        wb.emplace_back(new std::barrier<>(32));
        emu_warp_barrier[wi] = wb.back().get();
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < block; ++t)
        pool.emplace_back([=, &bar] {
            threadIdx.x = (unsigned)t;
            for (int b = 0; b < grid; ++b) {
                blockIdx.x = (unsigned)b;
                cta(p);
                bar.arrive_and_wait();  // CTAs share the one shared-memory buffer: run them in turn
            }
        });
    for (auto &th : pool) th.join();
    for (auto &q : emu_warp_barrier) q = nullptr;
    return 0;
}
#endif  // MPDATA_EMUL_NO_MARCH
