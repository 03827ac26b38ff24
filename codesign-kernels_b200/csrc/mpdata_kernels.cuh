// Device source of libmpdata_b200.so: the MMF MPDATA tracer step
// (mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90:477-640, cited as A:<line>) as ONE launch.
//
// The reference's OpenACC forms are 9 and 16 kernels that stream f, u, w and four scratch fields
// (uuu, www, mx, mn) through device memory between them (A:112-227, A:287-457).  Every slice is
// independent and one slice's seven x-z planes are (nx+6)*(nz-1) doubles each -- 115 KB at the
// reference sizes -- so here a CTA owns one slice, keeps all seven planes in shared memory for the
// whole step, and touches HBM once per input value and once per output value:
//   P0 load f,u,w (+ rho, rhow, adz columns)            P4 limiter ratios mx, mn
//   P1 first-order fluxes uuu, www; local extrema       P5 limited fluxes
//   P2 first-order update of f; flux = sum_i www        P6 final f (clipped at 0), flux += sum_i www
//   P3 antidiffusive fluxes; extrema of the updated f
// with one barrier between phases.  Sums over i are sequential per level, as the reference's are,
// and the file is compiled with -fmad=false, so the result is BIT-IDENTICAL to the CPU order.
//
// This header is also compiled for the host by tests/driver/mpdata_emul.cpp (test infrastructure).
#pragma once

namespace mpd_dev {

struct Problem {
    int ns, nx, nz;
    double *f;                                  // (ns, -2:nx+3, nzm)  inout
    const double *u, *w, *rho, *rhow, *adz;     // (ns, -1:nx+3, nzm) (ns, -1:nx+2, nz) (ns,nzm) (ns,nz) (ns,nzm)
    double *flux;                               // (ns, nz)  out, level nz untouched (A:482)
};

// doubles of shared memory one slice needs
__host__ __device__ inline size_t slice_doubles(int nx, int nz) {
    const size_t nzm = nz - 1;
    return (size_t)(nx + 6) * nzm + 2 * (size_t)(nx + 5) * nzm + 2 * (size_t)(nx + 4) * nz + 2 * (size_t)(nx + 2) * nzm +
           6 * (size_t)nz;
}

// Fortran MAX/MIN as compare-select, in the oracle's argument order (so even the sign of a zero
// agrees).  CUDA's fmax/fmin cost ~8 instructions each on sm_100a (DSETP pair, two selects, the
// NaN-quieting LOP3s and register moves); these are DSETP + two selects.
__device__ __forceinline__ double dmax(double a, double b) { return a > b ? a : b; }
__device__ __forceinline__ double dmin(double a, double b) { return a < b ? a : b; }
// statement functions A:500-503 (left-to-right products, as Fortran evaluates them)
__device__ __forceinline__ double andiff(double x1, double x2, double a, double b) { return (fabs(a) - a * a * b) * 0.5 * (x2 - x1); }
__device__ __forceinline__ double across(double x1, double a1, double a2) { return 0.03125 * a1 * a2 * x1; }
__device__ __forceinline__ double pp(double y) { return dmax(0.0, y); }
__device__ __forceinline__ double pn(double y) { return -dmin(0.0, y); }

// q -> (k = q / W, i = q % W) without an integer division: exact for q < 65536, W < 16384
// (checked exhaustively in tests/test_zzz_mpdata.py); a plane never has more than 29,056 points
// (227 KB of shared memory), so every index this kernel splits is in range.
__device__ __forceinline__ void split(int q, int W, float invW, int &k, int &i) {
    k = (int)(((float)q + 0.5f) * invW);
    i = q - k * W;
}

__global__ void __launch_bounds__(1024, 1) advect_kernel(Problem p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nx = p.nx, nz = p.nz, nzm = nz - 1, ns = p.ns;
    const int FX = nx + 6, UX = nx + 5, WX = nx + 4, MW = nx + 2;
    const int nxp1 = nx + 1, nxp2 = nx + 2;
    const int tid = threadIdx.x, nth = blockDim.x;
    const double eps = (double)1.e-10f;  // default-real literal assigned to a double (A:509)
    const float invUX = 1.0f / (float)UX, invWX = 1.0f / (float)WX, invMW = 1.0f / (float)MW;
    const float invP3 = 1.0f / (float)(nx + 3), invP5 = 1.0f / (float)nxp1;

    double *sf = reinterpret_cast<double *>(smem_raw);  // f   [k-1][i+2]
    double *su = sf + FX * nzm;                         // u   [k-1][i+1]
    double *sw = su + UX * nzm;                         // w   [k-1][i+1], nz levels
    double *suuu = sw + WX * nz;                        // uuu [k-1][i+1]
    double *swww = suuu + UX * nzm;                     // www [k-1][i+1], nz levels
    double *smx = swww + WX * nz;                       // mx  [k-1][i]
    double *smn = smx + MW * nzm;                       // mn  [k-1][i]
    double *srho = smn + MW * nzm;                      // per-level columns, nz each
    double *sdd = srho + nz, *sirho = sdd + nz, *siadz = sirho + nz, *sirhow = siadz + nz, *sflux = sirhow + nz;

#define SF(i, k) sf[((i) + 2) + FX * ((k)-1)]
#define SU(i, k) su[((i) + 1) + UX * ((k)-1)]
#define SW(i, k) sw[((i) + 1) + WX * ((k)-1)]
#define SUUU(i, k) suuu[((i) + 1) + UX * ((k)-1)]
#define SWWW(i, k) swww[((i) + 1) + WX * ((k)-1)]
#define SMX(i, k) smx[(i) + MW * ((k)-1)]
#define SMN(i, k) smn[(i) + MW * ((k)-1)]

    for (int sl = blockIdx.x; sl < ns; sl += gridDim.x) {
        // ---- P0: the slice's planes; plane index == shared index, slice stride in global memory
        for (int q = tid; q < FX * nzm; q += nth) sf[q] = p.f[sl + (size_t)ns * q];
        for (int q = tid; q < UX * nzm; q += nth) su[q] = p.u[sl + (size_t)ns * q];
        for (int q = tid; q < WX * nz; q += nth) sw[q] = p.w[sl + (size_t)ns * q];
        for (int k = tid + 1; k <= nzm; k += nth) {
            const double r = p.rho[sl + (size_t)ns * (k - 1)], a = p.adz[sl + (size_t)ns * (k - 1)];
            const int kb = k > 1 ? k - 1 : 1, kc = k < nzm ? k + 1 : nzm;
            srho[k - 1] = r;
            sdd[k - 1] = 2.0 / (double)(kc - kb) / a;                           // A:569, the same for every x
            sirho[k - 1] = 1.0 / r;                                             // A:552
            siadz[k - 1] = 1.0 / a;                                             // A:553
            sirhow[k - 1] = 1.0 / (p.rhow[sl + (size_t)ns * (k - 1)] * a);      // A:565
        }
        __syncthreads();

        // ---- P1: first-order (upwind) fluxes and the local extrema of f (A:511-541)
        for (int q = tid; q < UX * nzm; q += nth) {
            int k, i;
            split(q, UX, invUX, k, i);
            k += 1, i -= 1;  // i = -1..nx+3
            const double uu = SU(i, k);
            SUUU(i, k) = dmax(0.0, uu) * SF(i - 1, k) + dmin(0.0, uu) * SF(i, k);
            if (i <= nxp2) {
                const int kb = k > 1 ? k - 1 : 1;
                const double ww = SW(i, k);
                SWWW(i, k) = dmax(0.0, ww) * SF(i, kb) + dmin(0.0, ww) * SF(i, k);
                if (k == 1) SWWW(i, nz) = 0.0;
            }
            if (i >= 0 && i <= nxp1) {
                const int kb = k > 1 ? k - 1 : 1, kc = k < nzm ? k + 1 : nzm;
                const double a = SF(i - 1, k), b = SF(i + 1, k), c = SF(i, kb), d = SF(i, kc), e = SF(i, k);
                SMX(i, k) = dmax(dmax(dmax(dmax(a, b), c), d), e);
                SMN(i, k) = dmin(dmin(dmin(dmin(a, b), c), d), e);
            }
        }
        __syncthreads();

        // ---- P2: flux = sum_i www (sequential, A:542-547; the last threads take it, they own the
        //          fewest points of the update loop) and the first-order update of f (A:550-560)
        if (tid >= nth - nzm || nth < nzm) {
            for (int k = nth - tid; k <= nzm; k += nth) {
                double s = 0.0;
                for (int i = 1; i <= nx; ++i) s = s + SWWW(i, k);
                sflux[k - 1] = s;
            }
        }
        for (int q = tid; q < WX * nzm; q += nth) {
            int k, i;
            split(q, WX, invWX, k, i);
            k += 1, i -= 1;  // i = -1..nx+2
            SF(i, k) = SF(i, k) - (SUUU(i + 1, k) - SUUU(i, k) + (SWWW(i, k + 1) - SWWW(i, k)) * siadz[k - 1]) * sirho[k - 1];
        }
        __syncthreads();

        // ---- P3: antidiffusive fluxes from the updated f (A:561-586) and its extrema (A:589-600)
        for (int q = tid; q < (nx + 3) * nzm; q += nth) {
            int k, i;
            split(q, nx + 3, invP3, k, i);
            k += 1;  // i = 0..nx+2
            const int kb = k > 1 ? k - 1 : 1, kc = k < nzm ? k + 1 : nzm;
            const int ib = i - 1, ic = i + 1;
            const double dd = sdd[k - 1];
            SUUU(i, k) = andiff(SF(ib, k), SF(i, k), SU(i, k), sirho[k - 1]) -
                         across(dd * (SF(ib, kc) + SF(i, kc) - SF(ib, kb) - SF(i, kb)), SU(i, k),
                                SW(ib, k) + SW(ib, kc) + SW(i, k) + SW(i, kc)) * sirho[k - 1];
            if (i <= nxp1) {
                const double v = andiff(SF(i, kb), SF(i, k), SW(i, k), sirhow[k - 1]) -
                                 across(SF(ic, kb) + SF(ic, k) - SF(ib, kb) - SF(ib, k), SW(i, k),
                                        SU(i, kb) + SU(i, k) + SU(ic, k) + SU(ic, kb)) * sirho[k - 1];
                SWWW(i, k) = k == 1 ? 0.0 : v;  // www(:,:,:,1) = 0 (A:586)
                const double a = SF(ib, k), b = SF(ic, k), c = SF(i, kb), d = SF(i, kc), e = SF(i, k);
                SMX(i, k) = dmax(dmax(dmax(dmax(dmax(a, b), c), d), e), SMX(i, k));
                SMN(i, k) = dmin(dmin(dmin(dmin(dmin(a, b), c), d), e), SMN(i, k));
            }
        }
        if (tid < 2) SWWW(tid == 0 ? -1 : nxp2, 1) = 0.0;  // the two columns of level 1 the loop above does not visit
        __syncthreads();

        // ---- P4: limiter ratios (A:601-612), in place
        for (int q = tid; q < MW * nzm; q += nth) {
            int k, i;
            split(q, MW, invMW, k, i);
            k += 1;  // i = 0..nx+1
            const int kc = k < nzm ? k + 1 : nzm, ic = i + 1;
            const double fk = SF(i, k), r = srho[k - 1], ia = siadz[k - 1];
            const double ui = SUUU(i, k), uc = SUUU(ic, k), wk = SWWW(i, k), wc = SWWW(i, kc);
            SMX(i, k) = r * (SMX(i, k) - fk) / (pn(uc) + pp(ui) + ia * (pn(wc) + pp(wk)) + eps);
            SMN(i, k) = r * (fk - SMN(i, k)) / (pp(uc) + pn(ui) + ia * (pp(wc) + pn(wk)) + eps);
        }
        __syncthreads();

        // ---- P5: limited fluxes (A:613-625), in place (each reads only its own flux)
        for (int q = tid; q < nxp1 * nzm; q += nth) {
            int k, i;
            split(q, nxp1, invP5, k, i);
            k += 1, i += 1;  // i = 1..nx+1
            const int kb = k > 1 ? k - 1 : 1, ib = i - 1;
            const double uu = SUUU(i, k);
            SUUU(i, k) = pp(uu) * dmin(dmin(1.0, SMX(i, k)), SMN(ib, k)) - pn(uu) * dmin(dmin(1.0, SMX(ib, k)), SMN(i, k));
            if (i <= nx) {
                const double ww = SWWW(i, k);
                SWWW(i, k) = pp(ww) * dmin(dmin(1.0, SMX(i, k)), SMN(i, kb)) - pn(ww) * dmin(dmin(1.0, SMX(i, kb)), SMN(i, k));
            }
        }
        __syncthreads();

        // ---- P6: flux += sum_i www (A:623), final f clipped at zero (A:630-637), results to HBM
        if (tid >= nth - nzm || nth < nzm) {
            for (int k = nth - tid; k <= nzm; k += nth) {
                double s = sflux[k - 1];
                for (int i = 1; i <= nx; ++i) s = s + SWWW(i, k);
                p.flux[sl + (size_t)ns * (k - 1)] = s;
            }
        }
        for (int q = tid; q < WX * nzm; q += nth) {
            int k, i;
            split(q, WX, invWX, k, i);
            k += 1, i -= 1;  // i = -1..nx+2: every point of f the step changes
            double v = SF(i, k);
            if (i >= 1 && i <= nx)
                v = dmax(0.0, v - (SUUU(i + 1, k) - SUUU(i, k) + (SWWW(i, k + 1) - SWWW(i, k)) * siadz[k - 1]) * sirho[k - 1]);
            p.f[sl + (size_t)ns * (((i) + 2) + FX * (k - 1))] = v;
        }
        __syncthreads();  // the next slice of this CTA reuses the planes
    }
#undef SF
#undef SU
#undef SW
#undef SUUU
#undef SWWW
#undef SMX
#undef SMN
}

}  // namespace mpd_dev
