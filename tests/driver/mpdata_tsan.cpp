// TEST INFRASTRUCTURE ONLY -- runs the host emulation of the MPDATA device source (mpdata_emul.cpp) under
// ThreadSanitizer: CUDA threads are host threads and __syncthreads is a std::barrier there, so a shared-memory
// access pair that is not ordered by a barrier -- a missing __syncthreads in the kernel -- is reported as a data race.
// Built by tests/test_zzz_mpdata.py with -fsanitize=thread; MPDATA_EMUL_SOURCE may point at a modified copy of the
// emulator (the negative control removes one barrier and must be reported).
#ifndef MPDATA_EMUL_SOURCE
#define MPDATA_EMUL_SOURCE "mpdata_emul.cpp"
#endif
#include MPDATA_EMUL_SOURCE
#include <cstdio>
#include <random>
int main() {
    std::mt19937_64 g(1);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    int worst = 0;
    for (int chunks : {1, 3}) {
        const int ns = 3, nx = 12, nz = 20, nzm = nz - 1;
        std::vector<double> f((size_t)ns * (nx + 6) * nzm), u((size_t)ns * (nx + 5) * nzm), w((size_t)ns * (nx + 4) * nz), rho(ns * nzm), rhow(ns * nz), adz(ns * nzm), flux(ns * nz);
        for (auto *v : {&f, &u, &w, &rho, &rhow, &adz, &flux}) for (auto &x : *v) x = U(g);
        for (auto &x : u) x -= 0.5;
        for (auto &x : w) x -= 0.5;
        int rc = mpdata_emul_step(2, 64, chunks, ns, nx, nz, f.data(), u.data(), w.data(), rho.data(), rhow.data(), adz.data(), flux.data());
        printf("chunks %d rc %d f[100]=%g\n", chunks, rc, f[100]);
        worst |= rc;
    }
#ifndef MPDATA_EMUL_NO_MARCH
    {   // the marching form (mpdata_march.cuh): lanes are host threads, __syncwarp a barrier of the warp's 32
        const int ns = 4, nx = 7, nz = 12, nzm = nz - 1;
        std::vector<double> f((size_t)ns * (nx + 6) * nzm), u((size_t)ns * (nx + 5) * nzm), w((size_t)ns * (nx + 4) * nz), rho(ns * nzm), rhow(ns * nz), adz(ns * nzm), flux(ns * nz);
        for (auto *v : {&f, &u, &w, &rho, &rhow, &adz, &flux}) for (auto &x : *v) x = U(g);
        for (auto &x : u) x -= 0.5;
        for (auto &x : w) x -= 0.5;
        for (int spw : {1, 2}) {  // slices per warp
            std::vector<double> fc = f, fl = flux;
            int rc = mpdata_emul_march_step(2, 64, spw, ns, nx, nz, fc.data(), u.data(), w.data(), rho.data(), rhow.data(), adz.data(), fl.data());
            printf("march%d rc %d f[100]=%g\n", spw, rc, fc[100]);
            worst |= rc;
        }
    }
#endif
    return worst;
}
