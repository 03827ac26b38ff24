// TEST INFRASTRUCTURE ONLY -- runs the host emulation of the nested_loops device source (cke_emul.cpp) under
// ThreadSanitizer: CUDA threads are host threads and __syncthreads is a std::barrier there, so a shared-memory
// access pair not ordered by a barrier (a missing __syncthreads between staging the edge records and using them,
// or before the next block of edges overwrites them) is reported as a data race.  Built by
// tests/test_zz_nested.py with -fsanitize=thread; CKE_EMUL_SOURCE may point at a modified copy (negative control).
#ifndef CKE_EMUL_SOURCE
#define CKE_EMUL_SOURCE "cke_emul.cpp"
#endif
#include CKE_EMUL_SOURCE
#include <cstdio>
#include <random>
int main() {
    std::mt19937_64 g(3);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    const int nEdges = 45, nCells = 12, nLev = 10, nvldim = 12, nAdv = 6;
    std::vector<int> nadv(nEdges), kmin(nCells), kmax(nCells), cells((size_t)nEdges * nAdv);
    for (auto &x : nadv) x = 1 + (int)(U(g) * nAdv) % nAdv;
    for (int c = 0; c < nCells; ++c) { kmin[c] = 1 + (int)(U(g) * 3); kmax[c] = nLev - (int)(U(g) * 3); }
    for (auto &x : cells) x = 1 + (int)(U(g) * nCells) % nCells;
    std::vector<double> tracer((size_t)nCells * nvldim), mask((size_t)nCells * nvldim), ntf((size_t)nEdges * nvldim),
        hmask((size_t)nEdges * nvldim), c1((size_t)nEdges * nAdv), c3((size_t)nEdges * nAdv), out((size_t)nEdges * nvldim);
    for (auto *v : {&tracer, &mask, &ntf, &hmask, &c1, &c3}) for (auto &x : *v) x = U(g) - 0.3;
    int worst = 0;
    for (int impl : {1, 2})
        for (int unroll : {1, 5}) {
            int rc = cke_emul_iteration(unroll, impl, 3, 64, 4, nEdges, nCells, nLev, nvldim, nAdv, nadv.data(), kmin.data(), kmax.data(),
                                        cells.data(), tracer.data(), ntf.data(), hmask.data(), mask.data(), c1.data(), c3.data(), 2.14f,
                                        out.data());
            printf("impl %d unroll %d rc %d out[7]=%g\n", impl, unroll, rc, out[7]);
            worst |= rc;
        }
    return worst;
}
