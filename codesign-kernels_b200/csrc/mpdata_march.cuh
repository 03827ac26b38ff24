// Marching form of the MPDATA step (mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90:477-642, cited A:<line>): the
// same arithmetic as advect_kernel (mpdata_kernels.cuh), scheduled without a single CTA-wide barrier.
//
// ncu of advect_kernel at 9,472 slices (profiles/r3j_*): 41 % of the issue slots used; per issued instruction a warp waits
// 5.5 cycles at a barrier and 4.2 on the global loads of the phase that fills shared memory -- the seven barrier-separated
// phases themselves are the bound.  Here ONE WARP owns a slice and marches through its levels MG = 3 at a time; the six
// stages of the step run one behind the other on a sliding window of levels, and the window lives in rings of MR = 8 rows
// per field (slot of level k = k & 7) in the warp's private part of shared memory:
//
//   step t (B = 3 t)   stage                                               levels it works on
//   S1  first-order fluxes uuu, www; extrema of f (A:513-541)              B+1 .. B+3
//   S2  first-order update of f, in place; flux = sum_i www (A:542-560)    B   .. B+2     (needs www of the level above)
//   S3  antidiffusive fluxes from the updated f; its extrema (A:561-600)   B-1 .. B+1     (needs f of the level above)
//   S4  limiter ratios (A:601-612)                                         B-2 .. B        (needs www of the level above)
//   S5  limited fluxes (A:613-625)                                         B-2 .. B        (needs the ratios of the level below)
//   S6  final f, clipped at zero (A:630-637); flux += sum_i www; to HBM    B-3 .. B-1     (needs www of the level above)
//
// so level k is live from S1 (step ceil(k/3)-1) to S6 (four rows later): at most MG + 5 = 8 rows per field at a time, which
// is why the rings hold exactly 8.  Between stages there is only a __syncwarp(); the warps of a CTA never meet.  The next
// three levels of f, u, w and of the level columns are loaded into REGISTERS at the top of a step and stored into the rings
// at its end, into the slots S6 has just released -- the loads travel under the whole step.  A stage's points (3 levels x
// up to 38 columns) are spread over the lanes in 4 rounds; within a stage every round first loads and computes into
// registers, then all rounds store: the in-place updates read only their own point, so the rounds are independent and the
// compiler is free to interleave their dependency chains.  Sums over i stay sequential per level (one lane per level), the
// file is compiled with -fmad=false: BIT-IDENTICAL to the CPU order, as advect_kernel is.
//
// A stage's 3 x 33..38 points fill four rounds of 32 lanes to 77..89 % only.  With NS2 = 2 a warp marches through TWO slices
// at once (two sets of rings): 6 x 33..38 points fill seven (eight) rounds to 88..99 %, and there are twice as many
// independent rounds per stage to interleave.
//
// Shared memory: 8 x (7 nx + 34) doubles = 16.5 KB per slice in flight at nx = 32 -> 12 or 13 slices per SM.  Sizes are
// compile-time (the library instantiates the reference sizes A:6-8; any other size runs advect_kernel).
//
// This header is also compiled for the host by tests/driver/mpdata_emul.cpp (test infrastructure).
#pragma once

#include "mpdata_kernels.cuh"

namespace mpd_dev {

constexpr int MG = 3;  // levels a warp advances per step
constexpr int MR = 8;  // rows per ring
static_assert(MR == MG + 5 && (MR & (MR - 1)) == 0, "the rings hold MG + 5 rows and are indexed by k & (MR - 1)");

// doubles of shared memory one slice in flight needs (a warp: NS2 times that)
__host__ __device__ inline size_t march_warp_doubles(int nx) { return (size_t)MR * (7 * (size_t)nx + 28 + 6); }

#ifdef __CUDA_ARCH__
#define MPD_LDG(ptr) __ldg(ptr)
#else
#define MPD_LDG(ptr) (*(ptr))
#endif

template <int NX>
struct MarchRings {
    static constexpr int FX = NX + 6, UX = NX + 5, WX = NX + 4, MW = NX + 2;
    double *f, *u, *w, *uu, *ww, *mx, *mn;        // [MR][FX] [MR][UX] [MR][WX] [MR][UX] [MR][WX] [MR][MW] [MR][MW]
    double *rho, *irho, *iadz, *irhow, *dd, *fl;  // [MR] each: level columns and the first flux sum
    __device__ __forceinline__ explicit MarchRings(double *b) {
        f = b; u = f + MR * FX; w = u + MR * UX; uu = w + MR * WX; ww = uu + MR * UX; mx = ww + MR * WX; mn = mx + MR * MW;
        rho = mn + MR * MW; irho = rho + MR; iadz = irho + MR; irhow = iadz + MR; dd = irhow + MR; fl = dd + MR;
    }
    // the rings of the warp's slice number sub
    __device__ __forceinline__ MarchRings at(int sub) const { return MarchRings(f + sub * (MR * (7 * NX + 28 + 6))); }
};

// point pt of a stage whose rows are WIDTH wide -> (slice of the warp, row of the stage, column)
template <int NS2, int WIDTH>
__device__ __forceinline__ void mpoint(int pt, int &sub, int &row, int &col) {
    constexpr int NP = MG * WIDTH;
    sub = NS2 > 1 ? pt / NP : 0;
    const int q = pt - sub * NP;
    row = q / WIDTH;
    col = q - row * WIDTH;
}

__device__ __forceinline__ int mslot(int k) { return k & (MR - 1); }

// rounds of 32 lanes that cover n points
__host__ __device__ constexpr int mrounds(int n) { return (n + 31) / 32; }

// true when round rd of a stage with N points has a point for this lane and that point's level is a level of the slice.
// FULL: every level of the stage is (known per step), only the last round can be short.
template <bool FULL, int N>
__device__ __forceinline__ bool mactive(int rd, int pt, int k, int nzm) {
    return (rd + 1 < mrounds(N) || pt < N) && (FULL || (k >= 1 && k <= nzm));
}

// ---- S1: levels B+1 .. B+MG, columns i = -1 .. nx+3 -------------------------------------------------------------------
template <int NX, int NZ, int NS2, bool FULL>
__device__ __forceinline__ void march_s1(const MarchRings<NX> &R0, int B, int lane) {
    constexpr int FX = NX + 6, UX = NX + 5, WX = NX + 4, MW = NX + 2, nzm = NZ - 1, N = NS2 * MG * UX, NR = mrounds(N);
    double o_uu[NR], o_ww[NR], o_mx[NR], o_mn[NR];
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, ii;  // i = ii - 1
        const int pt = lane + 32 * rd;
        mpoint<NS2, UX>(pt, sub, row, ii);
        const int k = B + 1 + row;
        o_uu[rd] = o_ww[rd] = o_mx[rd] = o_mn[rd] = 0.0;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            const int s = mslot(k), sb = mslot(k > 1 ? k - 1 : 1), sc = mslot(k < nzm ? k + 1 : nzm);
            const double uv = R.u[s * UX + ii], fim = R.f[s * FX + ii], fi = R.f[s * FX + ii + 1];
            o_uu[rd] = dmax(0.0, uv) * fim + dmin(0.0, uv) * fi;
            // www (i <= nx+2) and the extrema (i = 0 .. nx+1) are computed for every column, from clamped columns where the
            // point has none, and stored for their own columns only: no branch inside a round, so that the rounds interleave
            const double wv = R.w[s * WX + (ii <= NX + 3 ? ii : NX + 3)], fkb = R.f[sb * FX + ii + 1];
            o_ww[rd] = dmax(0.0, wv) * fkb + dmin(0.0, wv) * fi;
            const double b = R.f[s * FX + (ii <= NX + 3 ? ii + 2 : NX + 5)], d = R.f[sc * FX + ii + 1];
            o_mx[rd] = dmax(dmax(dmax(dmax(fim, b), fkb), d), fi);
            o_mn[rd] = dmin(dmin(dmin(dmin(fim, b), fkb), d), fi);
        }
    }
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, ii;
        const int pt = lane + 32 * rd;
        mpoint<NS2, UX>(pt, sub, row, ii);
        const int k = B + 1 + row;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            const int s = mslot(k);
            R.uu[s * UX + ii] = o_uu[rd];
            if (ii <= NX + 3) R.ww[s * WX + ii] = o_ww[rd];
            if (ii >= 1 && ii <= NX + 2) { R.mx[s * MW + ii - 1] = o_mx[rd]; R.mn[s * MW + ii - 1] = o_mn[rd]; }
        }
    }
}

// ---- S2: levels B .. B+MG-1, columns i = -1 .. nx+2; and the first flux sum of the levels S1 has just finished -------
template <int NX, int NZ, int NS2, bool FULL>
__device__ __forceinline__ void march_s2(const MarchRings<NX> &R0, int B, int lane) {
    constexpr int FX = NX + 6, UX = NX + 5, WX = NX + 4, nzm = NZ - 1, N = NS2 * MG * WX, NR = mrounds(N);
    double o_f[NR];
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, ii;  // i = ii - 1
        const int pt = lane + 32 * rd;
        mpoint<NS2, WX>(pt, sub, row, ii);
        const int k = B + row;
        o_f[rd] = 0.0;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            const int s = mslot(k);
            const double wup = R.ww[mslot(k + 1) * WX + ii], wk1 = k < nzm ? wup : 0.0;  // www(:,:,:,nz) = 0 (A:511)
            o_f[rd] = R.f[s * FX + ii + 1] - (R.uu[s * UX + ii + 1] - R.uu[s * UX + ii] + (wk1 - R.ww[s * WX + ii]) * R.iadz[s]) * R.irho[s];
        }
    }
    {   // flux(k) = sum_i www(i,k), sequential as A:542-547 (www of these levels is not touched by this stage).  Lanes
        // 0 .. NS2 MG - 1 own a level of a slice each; the others walk the row of lane 0 along with them and drop the result,
        // so that the warp stays converged: lanes that skipped the loop would spin at the next __syncwarp() and take issue
        // slots while they wait (15 % of the executed instructions in the first ncu capture of this kernel, profiles/r3k_*).
        const int sub = lane / MG, k = B + 1 + lane % MG;
        const bool mine = lane < NS2 * MG && k <= nzm;
        const MarchRings<NX> R = R0.at(mine ? sub : 0);
        const double *wrow = R.ww + mslot(mine ? k : B + 1) * WX + 2;
        double s = 0.0;
        for (int i = 0; i < NX; ++i) s = s + wrow[i];
        if (mine) R.fl[mslot(k)] = s;
    }
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, ii;
        const int pt = lane + 32 * rd;
        mpoint<NS2, WX>(pt, sub, row, ii);
        const int k = B + row;
        if (mactive<FULL, N>(rd, pt, k, nzm)) R0.at(sub).f[mslot(k) * FX + ii + 1] = o_f[rd];
    }
}

// ---- S3: levels B-1 .. B+MG-2, columns i = 0 .. nx+2 ------------------------------------------------------------------
template <int NX, int NZ, int NS2, bool FULL>
__device__ __forceinline__ void march_s3(const MarchRings<NX> &R0, int B, int lane) {
    constexpr int FX = NX + 6, UX = NX + 5, WX = NX + 4, MW = NX + 2, nzm = NZ - 1, PW = NX + 3, N = NS2 * MG * PW, NR = mrounds(N);
    double o_uu[NR], o_ww[NR], o_mx[NR], o_mn[NR];
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, i;
        const int pt = lane + 32 * rd;
        mpoint<NS2, PW>(pt, sub, row, i);
        const int k = B - 1 + row;
        o_uu[rd] = o_ww[rd] = o_mx[rd] = o_mn[rd] = 0.0;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            const int s = mslot(k), sb = mslot(k > 1 ? k - 1 : 1), sc = mslot(k < nzm ? k + 1 : nzm);
            const double *fk = R.f + s * FX + i + 2, *fb = R.f + sb * FX + i + 2, *fc = R.f + sc * FX + i + 2;  // f(i, .)
            const double *uk = R.u + s * UX + i + 1, *ub = R.u + sb * UX + i + 1;
            const double *wk = R.w + s * WX + i + 1, *wc = R.w + sc * WX + i + 1;
            const double irho = R.irho[s];
            o_uu[rd] = andiff(fk[-1], fk[0], uk[0], irho) -
                       across(R.dd[s] * (fc[-1] + fc[0] - fb[-1] - fb[0]), uk[0], wk[-1] + wc[-1] + wk[0] + wc[0]) * irho;
            {   // i <= nx+1; computed for column nx+2 as well (every index stays inside its row) and not stored there
                // (the extrema of column nx+2 are not read from the rings: the slot a clamped column would name is another
                //  lane's, and that lane rewrites it in this stage)
                const bool has = i <= NX + 1;
                const double mx0 = has ? R.mx[s * MW + i] : 0.0, mn0 = has ? R.mn[s * MW + i] : 0.0;
                const double v = andiff(fb[0], fk[0], wk[0], R.irhow[s]) -
                                 across(fb[1] + fk[1] - fb[-1] - fk[-1], wk[0], ub[0] + uk[0] + uk[1] + ub[1]) * irho;
                o_ww[rd] = k == 1 ? 0.0 : v;  // www(:,:,:,1) = 0 (A:586)
                const double a = fk[-1], b = fk[1], cc = fb[0], d = fc[0], e = fk[0];
                o_mx[rd] = dmax(dmax(dmax(dmax(dmax(a, b), cc), d), e), mx0);
                o_mn[rd] = dmin(dmin(dmin(dmin(dmin(a, b), cc), d), e), mn0);
            }
        }
    }
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, i;
        const int pt = lane + 32 * rd;
        mpoint<NS2, PW>(pt, sub, row, i);
        const int k = B - 1 + row;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            const int s = mslot(k);
            R.uu[s * UX + i + 1] = o_uu[rd];
            if (i <= NX + 1) { R.ww[s * WX + i + 1] = o_ww[rd]; R.mx[s * MW + i] = o_mx[rd]; R.mn[s * MW + i] = o_mn[rd]; }
        }
    }
}

// ---- S4: levels B-2 .. B+MG-3, columns i = 0 .. nx+1: limiter ratios, in place ----------------------------------------
template <int NX, int NZ, int NS2, bool FULL>
__device__ __forceinline__ void march_s4(const MarchRings<NX> &R0, int B, int lane, double eps) {
    constexpr int FX = NX + 6, UX = NX + 5, WX = NX + 4, MW = NX + 2, nzm = NZ - 1, N = NS2 * MG * MW, NR = mrounds(N);
    double o_mx[NR], o_mn[NR];
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, i;
        const int pt = lane + 32 * rd;
        mpoint<NS2, MW>(pt, sub, row, i);
        const int k = B - 2 + row;
        o_mx[rd] = o_mn[rd] = 0.0;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            const int s = mslot(k), sc = mslot(k < nzm ? k + 1 : nzm);
            const double fk = R.f[s * FX + i + 2], r = R.rho[s], ia = R.iadz[s];
            const double ui = R.uu[s * UX + i + 1], uc = R.uu[s * UX + i + 2], wk = R.ww[s * WX + i + 1], wc = R.ww[sc * WX + i + 1];
            o_mx[rd] = div_ratio(r * (R.mx[s * MW + i] - fk), pn(uc) + pp(ui) + ia * (pn(wc) + pp(wk)) + eps);
            o_mn[rd] = div_ratio(r * (fk - R.mn[s * MW + i]), pp(uc) + pn(ui) + ia * (pp(wc) + pn(wk)) + eps);
        }
    }
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, i;
        const int pt = lane + 32 * rd;
        mpoint<NS2, MW>(pt, sub, row, i);
        const int k = B - 2 + row;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            R.mx[mslot(k) * MW + i] = o_mx[rd];
            R.mn[mslot(k) * MW + i] = o_mn[rd];
        }
    }
}

// ---- S5: levels B-2 .. B+MG-3, columns i = 1 .. nx+1: limited fluxes, in place ----------------------------------------
template <int NX, int NZ, int NS2, bool FULL>
__device__ __forceinline__ void march_s5(const MarchRings<NX> &R0, int B, int lane) {
    constexpr int UX = NX + 5, WX = NX + 4, MW = NX + 2, nzm = NZ - 1, PW = NX + 1, N = NS2 * MG * PW, NR = mrounds(N);
    double o_uu[NR], o_ww[NR];
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, i;
        const int pt = lane + 32 * rd;
        mpoint<NS2, PW>(pt, sub, row, i);
        i += 1;  // i = 1 .. nx+1
        const int k = B - 2 + row;
        o_uu[rd] = o_ww[rd] = 0.0;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            const int s = mslot(k), sb = mslot(k > 1 ? k - 1 : 1);
            const double mxi = R.mx[s * MW + i], mxb = R.mx[s * MW + i - 1], mni = R.mn[s * MW + i], mnb = R.mn[s * MW + i - 1];
            const double uv = R.uu[s * UX + i + 1];
            o_uu[rd] = pp(uv) * dmin(dmin(1.0, mxi), mnb) - pn(uv) * dmin(dmin(1.0, mxb), mni);
            {   // i <= nx; computed for column nx+1 as well and not stored there
                const double wv = R.ww[s * WX + i + 1];
                o_ww[rd] = pp(wv) * dmin(dmin(1.0, mxi), R.mn[sb * MW + i]) - pn(wv) * dmin(dmin(1.0, R.mx[sb * MW + i]), mni);
            }
        }
    }
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, i;
        const int pt = lane + 32 * rd;
        mpoint<NS2, PW>(pt, sub, row, i);
        i += 1;
        const int k = B - 2 + row;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            R.uu[mslot(k) * UX + i + 1] = o_uu[rd];
            if (i <= NX) R.ww[mslot(k) * WX + i + 1] = o_ww[rd];
        }
    }
}

// ---- S6: levels B-3 .. B+MG-4, columns i = -2 .. nx+3: results to HBM; and the second flux sum of the levels S5 has just
//      finished (B-2 .. B+MG-3).  sl: the warp's first slice; its slice number sub exists if sl + sub < ns ----------------
template <int NX, int NZ, int NS2, bool FULL>
__device__ __forceinline__ void march_s6(const MarchRings<NX> &R0, const Problem &p, int sl, int B, int lane) {
    constexpr int FX = NX + 6, UX = NX + 5, WX = NX + 4, nzm = NZ - 1, N = NS2 * MG * FX, NR = mrounds(N);
    const size_t ns = (size_t)p.ns;
    double o_f[NR];
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, ii;  // i = ii - 2
        const int pt = lane + 32 * rd;
        mpoint<NS2, FX>(pt, sub, row, ii);
        const int k = B - 3 + row;
        o_f[rd] = 0.0;
        if (mactive<FULL, N>(rd, pt, k, nzm)) {
            const MarchRings<NX> R = R0.at(sub);
            const int s = mslot(k);
            const double v = R.f[s * FX + ii];
            // i = 1 .. nx; the columns outside keep the first-order update / the input (they compute on clamped columns and drop it)
            const bool inner = ii >= 3 && ii <= NX + 2;
            const int ic = ii < 3 ? 3 : (ii > NX + 2 ? NX + 2 : ii);
            const double wup = R.ww[mslot(k + 1) * WX + ic - 1], wk1 = k < nzm ? wup : 0.0;
            const double vv = dmax(0.0, v - (R.uu[s * UX + ic] - R.uu[s * UX + ic - 1] + (wk1 - R.ww[s * WX + ic - 1]) * R.iadz[s]) * R.irho[s]);
            o_f[rd] = inner ? vv : v;
        }
    }
    {   // flux(k) += sum_i www(i,k), continuing the first sum (A:623); every lane walks a row (see march_s2)
        const int sub = lane / MG, k = B - 2 + lane % MG;
        const bool mine = lane < NS2 * MG && k >= 1 && k <= nzm;
        const int ks = mine ? k : (B >= 1 ? B : 1);  // a level whose www row and first sum exist
        const MarchRings<NX> R = R0.at(mine ? sub : 0);
        const double *wrow = R.ww + mslot(ks) * WX + 2;
        double s = R.fl[mslot(ks)];
        for (int i = 0; i < NX; ++i) s = s + wrow[i];
        if (mine && sl + sub < p.ns) p.flux[sl + sub + ns * (size_t)(k - 1)] = s;
    }
#pragma unroll
    for (int rd = 0; rd < NR; ++rd) {
        int sub, row, ii;
        const int pt = lane + 32 * rd;
        mpoint<NS2, FX>(pt, sub, row, ii);
        const int k = B - 3 + row;
        if (mactive<FULL, N>(rd, pt, k, nzm) && (NS2 == 1 || sl + sub < p.ns)) p.fout[sl + sub + ns * (size_t)(ii + FX * (k - 1))] = o_f[rd];
    }
}

// one stage of a step on levels lo .. hi: nothing if none of them is a level of the slice, the unguarded form if all are
#define MPD_MARCH_STAGE(fn, lo, hi, ...)                                        \
    do {                                                                        \
        if ((lo) >= 1 && (hi) <= nzm) fn<NX, NZ, NS2, true>(__VA_ARGS__);       \
        else if ((hi) >= 1 && (lo) <= nzm) fn<NX, NZ, NS2, false>(__VA_ARGS__); \
    } while (0)

// NS2 slices per warp (1 or 2).  Registers: 162 (NS2 = 1: 12 warps fill the register file) / up to 255 (NS2 = 2: 6 warps).
template <int NX, int NZ, int NS2>
__global__ void __launch_bounds__(NS2 == 1 ? 384 : 192, 1) advect_march_kernel(Problem p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int FX = NX + 6, UX = NX + 5, WX = NX + 4, nzm = NZ - 1;
    constexpr int PF = MG * FX, PU = MG * UX, PW = MG * WX;  // values of three levels of one slice
    constexpr int NF = NS2 * PF, NU = NS2 * PU, NW = NS2 * PW, NRF = mrounds(NF), NRU = mrounds(NU), NRW = mrounds(NW);
    constexpr int T = (nzm + 4 + MG - 1) / MG;  // steps until S6 has passed level nzm
    static_assert(NZ >= 3 && (NS2 == 1 || NS2 == 2) && 4 * MG * NS2 <= 32, "sizes");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const size_t ns = (size_t)p.ns;
    const double eps = (double)1.e-10f;  // default-real literal assigned to a double (A:509)
    const MarchRings<NX> R(reinterpret_cast<double *>(smem_raw) + (size_t)warp * (NS2 * MR * (7 * NX + 28 + 6)));
    // lanes 0 .. 4 MG NS2 - 1: one level column value each (irho, iadz, irhow, dd of a level of a slice)
    const int cj = lane % MG, cwhich = (lane / MG) & 3, csub = lane / (4 * MG);

    for (int sl = (blockIdx.x * wpb + warp) * NS2; sl < p.ns; sl += gridDim.x * wpb * NS2) {  // neighbouring warps: neighbouring slices
        double pf[NRF], pu[NRU], pw[NRW], pc1 = 1.0, pc2 = 1.0;
        // the slice a lane loads for: a warp whose last slice does not exist loads its first one twice (and stores nothing for it)
        auto lsl = [&](int sub) { return (size_t)(sl + sub < p.ns ? sl + sub : sl); };
        // levels kf .. kf+MG-1 of u, w and the level columns, kf+1 .. kf+MG of f: global -> registers
        auto fetch = [&](int kf) {
#pragma unroll
            for (int rd = 0; rd < NRF; ++rd) {
                const int pt = lane + 32 * rd, sub = NS2 > 1 ? pt / PF : 0, q = pt - sub * PF;
                pf[rd] = ((rd + 1 < NRF || pt < NF) && kf + 1 + q / FX <= nzm) ? MPD_LDG(p.f + lsl(sub) + ns * (size_t)(q + FX * kf)) : 0.0;
            }
#pragma unroll
            for (int rd = 0; rd < NRU; ++rd) {
                const int pt = lane + 32 * rd, sub = NS2 > 1 ? pt / PU : 0, q = pt - sub * PU;
                pu[rd] = ((rd + 1 < NRU || pt < NU) && kf + q / UX <= nzm) ? MPD_LDG(p.u + lsl(sub) + ns * (size_t)(q + UX * (kf - 1))) : 0.0;
            }
#pragma unroll
            for (int rd = 0; rd < NRW; ++rd) {
                const int pt = lane + 32 * rd, sub = NS2 > 1 ? pt / PW : 0, q = pt - sub * PW;
                pw[rd] = ((rd + 1 < NRW || pt < NW) && kf + q / WX <= nzm) ? MPD_LDG(p.w + lsl(sub) + ns * (size_t)(q + WX * (kf - 1))) : 0.0;
            }
            if (csub < NS2 && kf + cj <= nzm) {
                const size_t o = lsl(csub) + ns * (size_t)(kf + cj - 1);
                pc1 = MPD_LDG((cwhich == 0 ? p.rho : cwhich == 2 ? p.rhow : p.adz) + o);
                pc2 = cwhich == 2 ? MPD_LDG(p.adz + o) : 1.0;
            }
        };
        // registers -> rings
        auto commit = [&](int kf) {
#pragma unroll
            for (int rd = 0; rd < NRF; ++rd) {
                const int pt = lane + 32 * rd, sub = NS2 > 1 ? pt / PF : 0, q = pt - sub * PF, row = q / FX, k = kf + 1 + row;
                if ((rd + 1 < NRF || pt < NF) && k <= nzm) R.at(sub).f[mslot(k) * FX + q - row * FX] = pf[rd];
            }
#pragma unroll
            for (int rd = 0; rd < NRU; ++rd) {
                const int pt = lane + 32 * rd, sub = NS2 > 1 ? pt / PU : 0, q = pt - sub * PU, row = q / UX, k = kf + row;
                if ((rd + 1 < NRU || pt < NU) && k <= nzm) R.at(sub).u[mslot(k) * UX + q - row * UX] = pu[rd];
            }
#pragma unroll
            for (int rd = 0; rd < NRW; ++rd) {
                const int pt = lane + 32 * rd, sub = NS2 > 1 ? pt / PW : 0, q = pt - sub * PW, row = q / WX, k = kf + row;
                if ((rd + 1 < NRW || pt < NW) && k <= nzm) R.at(sub).w[mslot(k) * WX + q - row * WX] = pw[rd];
            }
            {   // the level columns: two divisions, executed by every lane (the idle ones divide 1 by 1) so that the warp stays converged
                const int k = kf + cj;
                const bool mine = csub < NS2 && k <= nzm;
                const int s = mslot(k), kbg = k > 1 ? k - 1 : 1, kcg = k < nzm ? k + 1 : nzm;
                const double num = 2.0 / (double)(mine && cwhich == 3 ? kcg - kbg : 2);  // dd = 2/(kc-kb)/adz (A:569); 2/2 = 1 for the other columns
                const double q = num / (pc1 * pc2);                                      // 1/rho (A:552), 1/adz (A:553), 1/(rhow*adz) (A:565)
                if (mine) {
                    const MarchRings<NX> Rs = R.at(csub);
                    if (cwhich == 0) { Rs.rho[s] = pc1; Rs.irho[s] = q; }
                    else if (cwhich == 1) Rs.iadz[s] = q;
                    else if (cwhich == 2) Rs.irhow[s] = q;
                    else Rs.dd[s] = q;
                }
            }
        };

        // levels 1 .. MG (f: 1 .. MG+1) before the first step
        fetch(1);
        for (int c = lane; c < NS2 * FX; c += 32) {
            const int sub = NS2 > 1 ? c / FX : 0;
            R.at(sub).f[mslot(1) * FX + c - sub * FX] = MPD_LDG(p.f + lsl(sub) + ns * (size_t)(c - sub * FX));
        }
        commit(1);
        __syncwarp();

        for (int t = 0; t < T; ++t) {
            const int B = MG * t, kf = B + MG + 1;
            if (kf <= nzm) fetch(kf);  // travels under the six stages
            MPD_MARCH_STAGE(march_s1, B + 1, B + MG, R, B, lane);
            __syncwarp();
            MPD_MARCH_STAGE(march_s2, B, B + MG - 1, R, B, lane);
            __syncwarp();
            MPD_MARCH_STAGE(march_s3, B - 1, B + MG - 2, R, B, lane);
            __syncwarp();
            MPD_MARCH_STAGE(march_s4, B - 2, B + MG - 3, R, B, lane, eps);
            __syncwarp();
            MPD_MARCH_STAGE(march_s5, B - 2, B + MG - 3, R, B, lane);
            __syncwarp();
            MPD_MARCH_STAGE(march_s6, B - 3, B + MG - 3, R, p, sl, B, lane);  // hi: the flux sum runs one level ahead of the f rows
            __syncwarp();
            if (kf <= nzm) commit(kf);
            __syncwarp();
        }
    }
}
#undef MPD_MARCH_STAGE

}  // namespace mpd_dev
