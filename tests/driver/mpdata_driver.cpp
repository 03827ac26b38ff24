// TEST DRIVER: the reference's `program test_advect`
// (mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90:2-66, cited as A:<line>) against the C ABI
// of libmpdata_b200.so, printing the lines the reference prints.  Arrays are laid out exactly as
// the Fortran program declares them (A:31-36: slice index fastest, x from -2 / -1, singleton y).
// Links the product library and, for the CPU leg only, the C oracle (test infrastructure).
//
//   mpdata_driver [nslices nx nz]      default 48 32 58 (A:6-8)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "mpdata_b200.h"

extern "C" int mpdo_advect_scalar2D_cpu(int ns, int nx, int nz, double *f, const double *u, const double *w, const double *rho,
                                        const double *rhow, const double *adz, double *flux);

struct Fields {
    std::vector<double> adz, f, u, w, rho, rhow, flux;
};

// init() of A:645-665: uniform [0,1) numbers in the reference's draw order.  The reference seeds the
// compiler's generator with 100; that stream cannot be reproduced, an LCG with the same seed stands in.
static void init(Fields &a, int ns, int nx, int nz) {
    const size_t nzm = nz - 1;
    unsigned long long s = 100;
    auto draw = [&s]() {
        s = (1103515245ULL * s + 12345ULL) % (1ULL << 31);
        return (double)(s >> 8) / (double)(1 << 23);
    };
    auto fill = [&](std::vector<double> &v, size_t n) {
        v.resize(n);
        for (auto &x : v) x = draw();
    };
    fill(a.adz, ns * nzm);
    fill(a.f, (size_t)ns * (nx + 6) * nzm);
    fill(a.u, (size_t)ns * (nx + 5) * nzm);
    fill(a.w, (size_t)ns * (nx + 4) * nz);
    fill(a.rho, ns * nzm);
    fill(a.rhow, (size_t)ns * nz);
    fill(a.flux, (size_t)ns * nz);
}

static double rel_l1(const std::vector<double> &a, const std::vector<double> &b) {  // compare(), A:679-683
    double num = 0.0, den = 0.0;
    for (size_t j = 0; j < a.size(); ++j) {
        num += std::fabs(a[j] - b[j]);
        den += std::fabs(b[j]);
    }
    return num / den;
}

int main(int argc, char **argv) {
    const int ns = argc > 3 ? atoi(argv[1]) : 48, nx = argc > 3 ? atoi(argv[2]) : 32, nz = argc > 3 ? atoi(argv[3]) : 58;
    Fields cpu, gpu;
    init(cpu, ns, nx, nz);  // call init(); call advect_scalar2D_cpu(...); call save()   (A:49-51)
    auto t0 = std::chrono::steady_clock::now();
    if (mpdo_advect_scalar2D_cpu(ns, nx, nz, cpu.f.data(), cpu.u.data(), cpu.w.data(), cpu.rho.data(), cpu.rhow.data(),
                                 cpu.adz.data(), cpu.flux.data())) {
        fprintf(stderr, "oracle: bad sizes\n");
        return 2;
    }
    printf(" CPU Timing:  %.16E\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());

    init(gpu, ns, nx, nz);  // call init(); call advect_scalar2D_<impl>(...); call compare()   (A:53-55)
    t0 = std::chrono::steady_clock::now();
    int rc = mpdata_advect_scalar2D(ns, nx, nz, gpu.f.data(), gpu.u.data(), gpu.w.data(), gpu.rho.data(), gpu.rhow.data(),
                                    gpu.adz.data(), gpu.flux.data());
    if (rc) {
        fprintf(stderr, "mpdata_advect_scalar2D failed (%d): %s\n", rc, mpdata_last_error());
        return 1;
    }
    printf(" B200 Timing (first call, allocation and transfers inside):  %.16E\n",
           std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    float ms = 0.f;  // the reference brackets the kernels only (A:110, A:238): CUDA events around the launch
    if ((rc = mpdata_time_steps(20, 0, &ms))) {
        fprintf(stderr, "mpdata_time_steps failed (%d): %s\n", rc, mpdata_last_error());
        return 1;
    }
    printf(" B200 Timing:  %.16E\n", ms * 1e-3);
    printf(" Relative L1 Error - f    :  %.16E\n", rel_l1(gpu.f, cpu.f));
    printf(" Relative L1 Error - flux :  %.16E\n", rel_l1(gpu.flux, cpu.flux));
    mpdata_finalize();
    return 0;
}
