// TEST INFRASTRUCTURE ONLY -- compiles the device source of libmpdata_b200.so
// (codesign-kernels_b200/csrc/mpdata_kernels.cuh, unmodified) for the HOST: the CUDA vocabulary is
// mapped onto blockDim.x host threads and a barrier, CTAs run one after another.  This is how the
// kernel's indexing, phase ordering (barrier-phased in-place updates in shared memory) and
// operation order are checked against the oracle on a box without a GPU; it is not a CPU path of
// the product (nothing in the package builds or loads it).
#include <barrier>
#include <cmath>
#include <cstddef>
#include <thread>
#include <vector>

struct emu_dim3 {
    unsigned x = 1, y = 1, z = 1;
};
static thread_local emu_dim3 threadIdx, blockIdx;
static emu_dim3 blockDim, gridDim;
static std::barrier<> *emu_barrier = nullptr;
static inline void __syncthreads() { emu_barrier->arrive_and_wait(); }

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
