// Device source of libmpdata_b200.so: the MMF MPDATA tracer step
// (mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90:477-642, cited as A:<line>) as ONE launch.
//
// The reference's OpenACC forms are 9 and 16 kernels that stream f, u, w and four scratch fields
// (uuu, www, mx, mn) through device memory between them (A:112-227, A:287-457).  Every slice is
// independent and one slice's seven x-z planes are (nx+6)*(nz-1) doubles each -- 115 KB at the
// reference sizes -- so here a CTA keeps all seven planes of its work item in shared memory for the
// whole step and touches HBM once per input value and once per output value:
//   P0 load f,u,w (+ rho, rhow, adz columns)            P4 limiter ratios mx, mn
//   P1 first-order fluxes uuu, www; local extrema       P5 limited fluxes
//   P2 first-order update of f; flux = sum_i www        P6 final f (clipped at 0), flux += sum_i www
//   P3 antidiffusive fluxes; extrema of the updated f
// with one barrier between phases.  A work item is a slice, or -- when there are fewer slices than
// SMs, as at the reference sizes (48 slices, 148 SMs) -- one of `chunks` level ranges of a slice:
// the final f(k) depends on the input within k-3..k+3 only, so a chunk loads its owned levels plus a
// halo of 3, recomputes the halo rows redundantly (their outermost values are wrong and never
// reach an owned output) and writes its owned levels.  Sums over i are sequential per level, as the
// reference's are, and the file is compiled with -fmad=false, so the result is BIT-IDENTICAL to the
// CPU order for every chunk count.
//
// This header is also compiled for the host by tests/driver/mpdata_emul.cpp (test infrastructure).
#pragma once

namespace mpd_dev {

struct Problem {
    int ns, nx, nz;
    const double *f;                            // (ns, -2:nx+3, nzm)  in
    double *fout;                               // same shape, the advanced field: a DIFFERENT buffer whenever a slice is
                                                // split into level chunks (a chunk's halo must see the other chunks' input)
    const double *u, *w, *rho, *rhow, *adz;     // (ns, -1:nx+3, nzm) (ns, -1:nx+2, nz) (ns,nzm) (ns,nz) (ns,nzm)
    double *flux;                               // (ns, nz)  out, level nz untouched (A:482)
};

constexpr int kHalo = 3;  // levels of input either side that the final f(k) and flux(k) depend on

// owned levels a..b (1-based, inclusive) of chunk c, and the levels klo..khi it keeps in shared memory
__host__ __device__ inline void chunk_range(int nzm, int chunks, int c, int &a, int &b, int &klo, int &khi) {
    a = 1 + (int)(((long long)c * nzm) / chunks);
    b = (int)(((long long)(c + 1) * nzm) / chunks);
    klo = a - kHalo > 1 ? a - kHalo : 1;
    khi = b + kHalo < nzm ? b + kHalo : nzm;
}
// most levels any chunk keeps
__host__ __device__ inline int max_local_levels(int nzm, int chunks) {
    int m = 0;
    for (int c = 0; c < chunks; ++c) {
        int a, b, klo, khi;
        chunk_range(nzm, chunks, c, a, b, klo, khi);
        if (khi - klo + 1 > m) m = khi - klo + 1;
    }
    return m;
}
// doubles of shared memory a work item of nl local levels needs
__host__ __device__ inline size_t item_doubles(int nx, int nl) {
    return (size_t)(nx + 6) * nl + 2 * (size_t)(nx + 5) * nl + (size_t)(nx + 4) * (2 * nl + 1) + 2 * (size_t)(nx + 2) * nl +
           6 * (size_t)nl;
}

// extra doubles of shared memory the prefetching form needs: a second set of the three input planes and two sets of the raw
// rho / adz / rhow columns
__host__ __device__ inline size_t prefetch_doubles(int nx, int nl) {
    return (size_t)(nx + 6) * nl + (size_t)(nx + 5) * nl + (size_t)(nx + 4) * nl + 6 * (size_t)nl;
}
// 8-byte asynchronous global -> shared copy (cp.async); the host emulation copies at once
__device__ __forceinline__ void async_copy8(double *dst, const double *src) {
#ifdef __CUDA_ARCH__
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
#else
    *dst = *src;
#endif
}
__device__ __forceinline__ void async_wait_all() {
#ifdef __CUDA_ARCH__
    asm volatile("cp.async.wait_all;" ::: "memory");
#endif
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
// num / den of the limiter ratios (A:601-612), bit-identical to the IEEE quotient for every input.  den is a sum of
// non-negative terms plus eps, and the numerator r * (mx - f) is exactly zero wherever f is a local extremum -- a zero
// numerator sends the lane through the out-of-line slow path of CUDA's double division (ncu, profiles/r3l_*: a call per
// division for ~6 lanes of every warp, 15 % of all executed instructions).  Those lanes divide 1 by den instead and take
// the zero IEEE gives them: (+-0) / den = +-0 for den > 0.
__device__ __forceinline__ double div_ratio(double num, double den) {
    const bool zero = num == 0.0 && den > 0.0;
    const double q = (zero ? 1.0 : num) / den;
    return zero ? num : q;
}

// q -> (k = q / W, i = q % W) without an integer division: exact for q < 65536, W < 16384
// (checked exhaustively in tests/test_zzz_mpdata.py); a plane never has more than 29,056 points
// (227 KB of shared memory), so every index this kernel splits is in range.
__device__ __forceinline__ void split(int q, int W, float invW, int &k, int &i) {
    k = (int)(((float)q + 0.5f) * invW);
    i = q - k * W;
}

// NXC, NZC > 0: the plane sizes as compile-time constants (the library instantiates the reference sizes, A:6-8: every plane
// stride, loop bound and index split below becomes an immediate); 0: taken from the problem at run time.
template <int NXC, int NZC>
__global__ void __launch_bounds__(1024, 1) advect_kernel(Problem p, int chunks, int nlmax, int prefetch) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nx = NXC > 0 ? NXC : p.nx, nz = NZC > 0 ? NZC : p.nz, nzm = nz - 1, ns = p.ns;
    const int FX = nx + 6, UX = nx + 5, WX = nx + 4, MW = nx + 2;
    const int nxp1 = nx + 1, nxp2 = nx + 2;
    const int tid = threadIdx.x, nth = blockDim.x;
    const double eps = (double)1.e-10f;  // default-real literal assigned to a double (A:509)
    const float invFX = 1.0f / (float)FX, invUX = 1.0f / (float)UX, invWX = 1.0f / (float)WX, invMW = 1.0f / (float)MW;
    const float invP3 = 1.0f / (float)(nx + 3), invP5 = 1.0f / (float)nxp1;

    // planes of nlmax local levels (www: one more, the level above the last), row = k - klo
    double *sfA = reinterpret_cast<double *>(smem_raw); // f   [row][i+2]
    double *suA = sfA + FX * nlmax;                     // u   [row][i+1]
    double *swA = suA + UX * nlmax;                     // w   [row][i+1]
    double *suuu = swA + WX * nlmax;                    // uuu [row][i+1]
    double *swww = suuu + UX * nlmax;                   // www [row][i+1], nlmax + 1 rows
    double *smx = swww + WX * (nlmax + 1);              // mx  [row][i]
    double *smn = smx + MW * nlmax;                     // mn  [row][i]
    double *srho = smn + MW * nlmax;                    // per-level columns, nlmax each
    double *sdd = srho + nlmax, *sirho = sdd + nlmax, *siadz = sirho + nlmax, *sirhow = siadz + nlmax, *sflux = sirhow + nlmax;
    // prefetching form (prefetch != 0, a CTA walks several items): a second set of the input planes and two sets of the raw
    // rho / adz / rhow columns; the loads of item n+1 travel (cp.async) while item n is computed
    double *sfB = sflux + nlmax, *suB = sfB + FX * nlmax, *swB = suB + UX * nlmax, *sraw = swB + WX * nlmax;  // sraw [set][rho,adz,rhow][nlmax]
    const int nitems = ns * chunks;
    auto fetch_item = [&](int item, int set) { // every thread: its share of the item's f, u, w planes and level columns
        const int c = item / ns, sl = item - c * ns;
        int ka, kz, klo, khi;
        chunk_range(nzm, chunks, c, ka, kz, klo, khi);
        const int nl = khi - klo + 1;
        double *df = set ? sfB : sfA, *du = set ? suB : suA, *dw = set ? swB : swA, *dr = sraw + set * 3 * nlmax;
        for (int q = tid; q < FX * nl; q += nth) async_copy8(df + q, p.f + sl + (size_t)ns * (q + FX * (klo - 1)));
        for (int q = tid; q < UX * nl; q += nth) async_copy8(du + q, p.u + sl + (size_t)ns * (q + UX * (klo - 1)));
        for (int q = tid; q < WX * nl; q += nth) async_copy8(dw + q, p.w + sl + (size_t)ns * (q + WX * (klo - 1)));
        for (int k = klo + tid; k <= khi; k += nth) {
            async_copy8(dr + (k - klo), p.rho + sl + (size_t)ns * (k - 1));
            async_copy8(dr + nlmax + (k - klo), p.adz + sl + (size_t)ns * (k - 1));
            async_copy8(dr + 2 * nlmax + (k - klo), p.rhow + sl + (size_t)ns * (k - 1));
        }
    };
    if (prefetch && (int)blockIdx.x < nitems) fetch_item(blockIdx.x, 0);
    int it = 0;

#define SF(i, k) sf[((i) + 2) + FX * ((k)-klo)]
#define SU(i, k) su[((i) + 1) + UX * ((k)-klo)]
#define SW(i, k) sw[((i) + 1) + WX * ((k)-klo)]
#define SUUU(i, k) suuu[((i) + 1) + UX * ((k)-klo)]
#define SWWW(i, k) swww[((i) + 1) + WX * ((k)-klo)]
#define SMX(i, k) smx[(i) + MW * ((k)-klo)]
#define SMN(i, k) smn[(i) + MW * ((k)-klo)]
#define COL(a, k) a[(k)-klo]
// neighbour levels: the reference's clamps at levels 1 and nzm (A:515-516); at an inner chunk edge the
// clamp keeps the index in range and the value it yields only feeds halo rows
#define KB(k) ((k) > klo ? (k)-1 : klo)
#define KC(k) ((k) < khi ? (k) + 1 : khi)

    for (int item = blockIdx.x; item < nitems; item += gridDim.x, ++it) {
        const int c = item / ns, sl = item - c * ns;  // neighbouring CTAs: neighbouring slices, same levels
        int ka, kz, klo, khi;
        chunk_range(nzm, chunks, c, ka, kz, klo, khi);
        const int nl = khi - klo + 1;
        const int set = prefetch ? (it & 1) : 0;
        double *sf = set ? sfB : sfA, *su = set ? suB : suA, *sw = set ? swB : swA;

        if (prefetch) {
            // ---- P0 (prefetching form): this item's planes were requested one item ago; every thread's copies have landed
            //      after its wait, everybody's after the barrier -- which is also the barrier that ends the previous item
            async_wait_all();
            __syncthreads();
            if (item + (int)gridDim.x < nitems) fetch_item(item + gridDim.x, set ^ 1); // travels while this item is computed
            const double *raw = sraw + set * 3 * nlmax;
            for (int k = klo + tid; k <= khi; k += nth) {
                const double r = raw[k - klo], a = raw[nlmax + (k - klo)];
                const int kbg = k > 1 ? k - 1 : 1, kcg = k < nzm ? k + 1 : nzm;
                COL(srho, k) = r;
                COL(sdd, k) = 2.0 / (double)(kcg - kbg) / a;
                COL(sirho, k) = 1.0 / r;
                COL(siadz, k) = 1.0 / a;
                COL(sirhow, k) = 1.0 / (raw[2 * nlmax + (k - klo)] * a);
            }
        } else {
            // ---- P0: the item's planes; slice stride in global memory
            for (int q = tid; q < FX * nl; q += nth) sf[q] = p.f[sl + (size_t)ns * (q + FX * (klo - 1))];
            for (int q = tid; q < UX * nl; q += nth) su[q] = p.u[sl + (size_t)ns * (q + UX * (klo - 1))];
            for (int q = tid; q < WX * nl; q += nth) sw[q] = p.w[sl + (size_t)ns * (q + WX * (klo - 1))];
            for (int k = klo + tid; k <= khi; k += nth) {
                const double r = p.rho[sl + (size_t)ns * (k - 1)], a = p.adz[sl + (size_t)ns * (k - 1)];
                const int kbg = k > 1 ? k - 1 : 1, kcg = k < nzm ? k + 1 : nzm;
                COL(srho, k) = r;
                COL(sdd, k) = 2.0 / (double)(kcg - kbg) / a;                          // A:569, the same for every x
                COL(sirho, k) = 1.0 / r;                                              // A:552
                COL(siadz, k) = 1.0 / a;                                              // A:553
                COL(sirhow, k) = 1.0 / (p.rhow[sl + (size_t)ns * (k - 1)] * a);       // A:565
            }
        }
        __syncthreads();

        // ---- P1: first-order (upwind) fluxes and the local extrema of f (A:511-541)
        for (int q = tid; q < UX * nl; q += nth) {
            int k, i;
            split(q, UX, invUX, k, i);
            k += klo, i -= 1;  // i = -1..nx+3
            const double uu = SU(i, k);
            SUUU(i, k) = dmax(0.0, uu) * SF(i - 1, k) + dmin(0.0, uu) * SF(i, k);
            if (i <= nxp2) {
                const double ww = SW(i, k);
                SWWW(i, k) = dmax(0.0, ww) * SF(i, KB(k)) + dmin(0.0, ww) * SF(i, k);
                if (k == klo) SWWW(i, khi + 1) = 0.0;  // www(:,:,:,nz) = 0 (A:511) when khi = nzm; unused otherwise
            }
            if (i >= 0 && i <= nxp1) {
                const double a = SF(i - 1, k), b = SF(i + 1, k), cc = SF(i, KB(k)), d = SF(i, KC(k)), e = SF(i, k);
                SMX(i, k) = dmax(dmax(dmax(dmax(a, b), cc), d), e);
                SMN(i, k) = dmin(dmin(dmin(dmin(a, b), cc), d), e);
            }
        }
        __syncthreads();

        // ---- P2: flux = sum_i www (sequential, A:542-547; the last threads take it, they own the
        //          fewest points of the update loop) and the first-order update of f (A:550-560)
        if (tid >= nth - nl || nth < nl) {
            for (int r = nth - tid; r <= nl; r += nth) {
                const int k = klo + r - 1;
                double s = 0.0;
                for (int i = 1; i <= nx; ++i) s = s + SWWW(i, k);
                COL(sflux, k) = s;
            }
        }
        for (int q = tid; q < WX * nl; q += nth) {
            int k, i;
            split(q, WX, invWX, k, i);
            k += klo, i -= 1;  // i = -1..nx+2
            SF(i, k) = SF(i, k) - (SUUU(i + 1, k) - SUUU(i, k) + (SWWW(i, k + 1) - SWWW(i, k)) * COL(siadz, k)) * COL(sirho, k);
        }
        __syncthreads();

        // ---- P3: antidiffusive fluxes from the updated f (A:561-586) and its extrema (A:589-600)
        for (int q = tid; q < (nx + 3) * nl; q += nth) {
            int k, i;
            split(q, nx + 3, invP3, k, i);
            k += klo;  // i = 0..nx+2
            const int kb = KB(k), kc = KC(k);
            const int ib = i - 1, ic = i + 1;
            const double dd = COL(sdd, k);
            SUUU(i, k) = andiff(SF(ib, k), SF(i, k), SU(i, k), COL(sirho, k)) -
                         across(dd * (SF(ib, kc) + SF(i, kc) - SF(ib, kb) - SF(i, kb)), SU(i, k),
                                SW(ib, k) + SW(ib, kc) + SW(i, k) + SW(i, kc)) * COL(sirho, k);
            if (i <= nxp1) {
                const double v = andiff(SF(i, kb), SF(i, k), SW(i, k), COL(sirhow, k)) -
                                 across(SF(ic, kb) + SF(ic, k) - SF(ib, kb) - SF(ib, k), SW(i, k),
                                        SU(i, kb) + SU(i, k) + SU(ic, k) + SU(ic, kb)) * COL(sirho, k);
                SWWW(i, k) = k == 1 ? 0.0 : v;  // www(:,:,:,1) = 0 (A:586)
                const double a = SF(ib, k), b = SF(ic, k), cc = SF(i, kb), d = SF(i, kc), e = SF(i, k);
                SMX(i, k) = dmax(dmax(dmax(dmax(dmax(a, b), cc), d), e), SMX(i, k));
                SMN(i, k) = dmin(dmin(dmin(dmin(dmin(a, b), cc), d), e), SMN(i, k));
            }
        }
        if (tid < 2 && klo == 1) SWWW(tid == 0 ? -1 : nxp2, 1) = 0.0;  // the two columns of level 1 the loop above does not visit
        __syncthreads();

        // ---- P4: limiter ratios (A:601-612), in place
        for (int q = tid; q < MW * nl; q += nth) {
            int k, i;
            split(q, MW, invMW, k, i);
            k += klo;  // i = 0..nx+1
            const int kc = KC(k), ic = i + 1;
            const double fk = SF(i, k), r = COL(srho, k), ia = COL(siadz, k);
            const double ui = SUUU(i, k), uc = SUUU(ic, k), wk = SWWW(i, k), wc = SWWW(i, kc);
            SMX(i, k) = div_ratio(r * (SMX(i, k) - fk), pn(uc) + pp(ui) + ia * (pn(wc) + pp(wk)) + eps);
            SMN(i, k) = div_ratio(r * (fk - SMN(i, k)), pp(uc) + pn(ui) + ia * (pp(wc) + pn(wk)) + eps);
        }
        __syncthreads();

        // ---- P5: limited fluxes (A:613-625), in place (each reads only its own flux)
        for (int q = tid; q < nxp1 * nl; q += nth) {
            int k, i;
            split(q, nxp1, invP5, k, i);
            k += klo, i += 1;  // i = 1..nx+1
            const int kb = KB(k), ib = i - 1;
            const double uu = SUUU(i, k);
            SUUU(i, k) = pp(uu) * dmin(dmin(1.0, SMX(i, k)), SMN(ib, k)) - pn(uu) * dmin(dmin(1.0, SMX(ib, k)), SMN(i, k));
            if (i <= nx) {
                const double ww = SWWW(i, k);
                SWWW(i, k) = pp(ww) * dmin(dmin(1.0, SMX(i, k)), SMN(i, kb)) - pn(ww) * dmin(dmin(1.0, SMX(i, kb)), SMN(i, k));
            }
        }
        __syncthreads();

        // ---- P6, owned levels ka..kz only: flux += sum_i www (A:623), final f clipped at zero (A:630-637), results to HBM
        const int no = kz - ka + 1;
        if (tid >= nth - no || nth < no) {
            for (int r = nth - tid; r <= no; r += nth) {
                const int k = ka + r - 1;
                double s = COL(sflux, k);
                for (int i = 1; i <= nx; ++i) s = s + SWWW(i, k);
                p.flux[sl + (size_t)ns * (k - 1)] = s;
            }
        }
        for (int q = tid; q < FX * no; q += nth) {
            int k, i;
            split(q, FX, invFX, k, i);
            k += ka, i -= 2;  // i = -2..nx+3: the step changes -1..nx+2, the two outer columns are carried over
            double v = SF(i, k);
            if (i >= 1 && i <= nx)
                v = dmax(0.0, v - (SUUU(i + 1, k) - SUUU(i, k) + (SWWW(i, k + 1) - SWWW(i, k)) * COL(siadz, k)) * COL(sirho, k));
            p.fout[sl + (size_t)ns * ((i + 2) + FX * (k - 1))] = v;
        }
        if (!prefetch) __syncthreads();  // the next item of this CTA reuses the planes (prefetching form: the barrier at the top of the next item)
    }
#undef SF
#undef SU
#undef SW
#undef SUUU
#undef SWWW
#undef SMX
#undef SMN
#undef COL
#undef KB
#undef KC
}

}  // namespace mpd_dev
