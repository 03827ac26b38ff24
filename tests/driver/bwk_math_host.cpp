// TEST INFRASTRUCTURE ONLY -- the ARITHMETIC of the hot kernel compiled for the host.
//
// codesign-kernels_b200/csrc/bwk_math.cuh holds what a thread of the sweep kernels computes on its slab (laplace_slab: the
// folded form of gradient_sphere -> tensorVisc -> divergence_sphere_wk, F90:109-182) and what fold_metric_kernel computes per
// GLL point (fold_metric_point).  bwk_kernels.cu includes that header for the device; this file includes THE SAME SOURCE
// for the host: __constant__ Dvv becomes a plain array, fma is <cmath>'s correctly rounded fma, the metric loader reads
// plain memory.  tests/test_kernel_math.py compares it with the oracle and with the translated reference on the CPU, so a
// change of the kernel's arithmetic is seen without a GPU.  Not the product path, never shipped; the data movement (TMA rings,
// mbarriers, strips, halo) is not emulated here -- the GPU tests cover that.
//
// Built by tests/driver/Makefile with -ffp-contract=off: only the kernel's explicit fma() calls are fused.  nvcc also contracts
// the plain a*b + c*d of fold_metric_point (-fmad=true), so this build and the GPU agree to rounding (<= 1e-15 relative on the
// metric), not to the bit; laplace_slab itself has no contractible expression left.
#include <cmath>
#include <cstdint>
#include <vector>

struct double2 {
    double x, y;
};
static double c_dvv[16];  // c_dvv[i + 4*l] == Dvv(i+1,l+1), as the device constant
using std::fma;

#include "../../codesign-kernels_b200/csrc/bwk_math.cuh"

namespace {
struct HostMetric {
    const double *m;
    double2 operator()(int pt, int half) const { return double2{m[4 * pt + 2 * half], m[4 * pt + 2 * half + 1]}; }
};
}  // namespace

extern "C" {

// fold_metric_kernel for nelemd elements: metric[nelemd][64]
int bwk_math_host_fold_metric(const double *elem, double *metric, int64_t nelemd, int use_tensor) {
    for (int64_t e = 0; e < nelemd; ++e)
        for (int pt = 0; pt < 16; ++pt) fold_metric_point(elem + e * 144, pt, use_tensor, metric + e * 64 + 4 * pt);
    return 0;
}

// what bwk_run(nets, nete) computes: qtens(np,np,nlev,qsize,nelemd) in place, elements nets..nete (1-based, inclusive)
int bwk_math_host_laplace(const double *elem, double *qtens, const double *dvv, int nlev, int qsize, int nelemd, int nets, int nete) {
    if (nets < 1 || nete > nelemd) return 1;
    for (int i = 0; i < 16; ++i) c_dvv[i] = dvv[i];
    std::vector<double> metric((size_t)64 * nelemd);
    bwk_math_host_fold_metric(elem, metric.data(), nelemd, 1);
    const int64_t kq = (int64_t)nlev * qsize;
    for (int64_t e = nets - 1; e < nete; ++e)
        for (int64_t s = 0; s < kq; ++s) {
            double v[16];
            double *slab = qtens + (e * kq + s) * 16;
            for (int t = 0; t < 16; ++t) v[t] = slab[t];
            laplace_slab(v, HostMetric{metric.data() + 64 * e});
            for (int t = 0; t < 16; ++t) slab[t] = v[t];
        }
    return 0;
}

}  // extern "C"
