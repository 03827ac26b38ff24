// bwk_math.cuh -- the ARITHMETIC of the hot path, separated from the data movement so that the very same source also compiles
// for the host: tests/driver/bwk_math_host.cpp includes this file with `c_dvv` as a plain array and `fma` from <cmath>, and the
// CPU suite checks it against the oracle and the translated reference (tests/test_kernel_math.py) -- a change of the kernel's
// arithmetic is seen without a GPU.  Included by bwk_kernels.cu inside namespace bwk, after `__constant__ double c_dvv[16]`.
#pragma once

#ifdef __CUDACC__
#define BWK_MATH_FN __device__ __forceinline__
#else
#define BWK_MATH_FN static inline
#endif

// F90:14: rrearth is a default-real literal -> float32-rounded value 0x1.65e9f8p-23
#define BWK_RREARTH 0x1.65e9f8p-23

// ------------------------------------------------------------------------------------------
// The operator on one register-resident slab.  M points at the element's folded metric,
// 4 doubles per GLL point: {Ma, Mb, Mc, Md} with w1 = Ma dx + Mb dy, w2 = Mc dx + Md dy.
// s[i + 4*j] = s(i+1,j+1).  Overwrites s with laplace_sphere_wk(s) (F90:164-182).
// ------------------------------------------------------------------------------------------
template <typename MetricLoader>
BWK_MATH_FN void laplace_slab(double (&s)[16], MetricLoader M)
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

// ------------------------------------------------------------------------------------------
// Folded metric of one GLL point: M = -(rrearth^2) * spheremp * Dinv * tensorVisc * Dinv^T, 4 doubles.
// E = the element's element_t record (F90:23-27 flattened: Dinv(i,j,c,a) at 16c+32a, spheremp at 64, tensorVisc at 80).
// ------------------------------------------------------------------------------------------
BWK_MATH_FN void fold_metric_point(const double *E, int pt, int use_tensor, double *M)
{
    const double D11 = E[pt], D21 = E[16 + pt], D12 = E[32 + pt], D22 = E[48 + pt]; // Dinv(i,j,c,a) at 16c+32a
    const double sp = E[64 + pt];
    // use_tensor = 0: laplace_sphere_wk with var_coef = .false. (upstream HOMME's first application under tensor hyperviscosity)
    const double T11 = use_tensor ? E[80 + pt] : 1.0, T21 = use_tensor ? E[96 + pt] : 0.0;
    const double T12 = use_tensor ? E[112 + pt] : 0.0, T22 = use_tensor ? E[128 + pt] : 1.0;
    // g = T * D^T * (dx,dy)  (F90:128-133 then F90:175-178)
    const double G1x = T11 * D11 + T12 * D12, G1y = T11 * D21 + T12 * D22;
    const double G2x = T21 * D11 + T22 * D12, G2y = T21 * D21 + T22 * D22;
    // vt = D * g  (F90:147-148), then * spheremp * rrearth (F90:155-156) and the gradient's rrearth (F90:124-125)
    const double f = -(BWK_RREARTH * BWK_RREARTH) * sp;
    M[0] = f * (D11 * G1x + D12 * G2x);
    M[1] = f * (D11 * G1y + D12 * G2y);
    M[2] = f * (D21 * G1x + D22 * G2x);
    M[3] = f * (D21 * G1y + D22 * G2y);
}
