// libmpdata_b200.so -- host side of the MMF MPDATA tracer step (include/mpdata_b200.h).
// sm_100a only, no CPU fallback: every entry point fails without a B200-class device.
// Compiled with -fmad=false (see mpdata_kernels.cuh: the result is bit-identical to the CPU order).
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/mpdata_b200.h"
#include "mpdata_kernels.cuh"
#include "mpdata_march.cuh"

namespace {

struct State {
    bool ready = false, uploaded = false;
    int ns = 0, nx = 0, nz = 0;
    size_t nf = 0, nu = 0, nw = 0, n2m = 0, n2 = 0;  // element counts of the arrays
    double *f = nullptr, *f1 = nullptr, *f0 = nullptr, *u = nullptr, *w = nullptr, *rho = nullptr, *rhow = nullptr, *adz = nullptr,
           *flux = nullptr, *flush = nullptr;
    size_t flush_bytes = 0;
    cudaStream_t stream = nullptr;
    int block = 0, chunks = 0;  // 0 = chosen by the library (see plan_launch)
    int prefetch = 0;           // the launch uses the prefetching form (see plan_launch)
    int force_generic = 0;      // MPDATA_B200_GENERIC=1: the run-time-size kernel even at the reference sizes (A/B, tests)
    int form = 0, march_warps = 0;  // mpdata_set_form: 0 = chosen by the library, 1 = barrier-phased CTA per slice, 2 / 3 = marching warp per slice / per two slices
    int use_form = 1;
    int use_block = 1024, use_chunks = 1, nlmax = 0, grid = 0, sms = 0, smem_optin = 0, smem_sm = 0;
    size_t smem = 0;
} g;

char g_err[512];
bool g_failed = false;
long long g_launches = 0;

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    g_failed = true;
    return code;
}

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) return fail(4, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

int release() {
    double **ptrs[] = {&g.f, &g.f1, &g.f0, &g.u, &g.w, &g.rho, &g.rhow, &g.adz, &g.flux, &g.flush};
    for (double **p : ptrs) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    if (g.stream) cudaStreamDestroy(g.stream);
    g.stream = nullptr;
    g.ready = g.uploaded = false;
    return 0;
}

// Work decomposition for the current sizes.  chunks: a slice's levels are split (with a halo of 3
// recomputed levels, mpdata_kernels.cuh) only while there are fewer slices than SMs, so that
// slices x chunks fills the SMs once (48 slices -> 3 chunks -> 144 CTAs on 148 SMs); block: about two
// points of the widest plane per thread.  Results do not depend on either.
// The marching forms (mpdata_march.cuh): a warp per slice (form 2) or per pair of slices (form 3), `warps` of them per CTA,
// as many CTAs per SM as their rings allow.
}  // namespace

// Which form is faster at the reference plane sizes (32 x 58), from the measured behaviour of the two kernels on a B200
// (profiles/r3n_*, r3o_*, r3p_*; times in ms):
//   barrier-phased, two half-slice CTAs per SM: linear in the slice count, 0.1325 us per slice on 148 SMs
//     (1,184 slices 0.159, 2,368 0.320, 4,736 0.627, 9,472 1.231);
//   marching, 6 warps per CTA, two CTAs per SM: a warp needs ~0.197 ms per slice whatever its neighbours do, so the time goes
//     in waves of 12 slices per SM (1,776 slices on 148 SMs): 1,776 slices 0.200, 3,552 0.393, 7,104 0.786, and a partial last
//     wave costs 0.12 (a third full), 0.16 (0.46) to 0.19 (two thirds): 2,368 0.328, 2,600 0.359, 4,736 0.584, 9,472 1.096,
//     1,184 alone 0.193, 1,600 alone 0.196.
// Host only; exported so that the CPU tests can check the choice against those measurements.
extern "C" int mpdata_march_wins(long long nslices, int sms) {
    if (nslices < 1 || sms < 1) return 0;
    const double wave = 12.0 * sms;                       // slices in flight: 2 CTAs x 6 warps per SM
    const double scale = 148.0 / sms;                     // the per-slice rate of the phased kernel was measured on 148 SMs
    const double phased = 0.1325e-3 * scale * (double)nslices;
    const long long full = (long long)((double)nslices / wave);
    const double r = ((double)nslices - (double)full * wave) / wave;  // fill of the last wave, 0..1
    double last = 0.0;
    if (r > 0.0)
        last = r <= 1.0 / 3 ? 0.08 + 0.04 * (3 * r) : r <= 0.5 ? 0.12 + 0.05 * (6 * r - 2) : r <= 2.0 / 3 ? 0.17 + 0.02 * (6 * r - 3) : 0.19 + 0.007 * (3 * r - 2);
    const double march = 0.197 * (double)full + last;
    return march < 0.97 * phased ? 1 : 0;  // within the model's error (2,600 slices: 0.359 against 0.350 ms) the phased kernel stays
}

namespace {

template <int NS2>
int plan_march(int default_warps = 0) {
    auto kernel = mpd_dev::advect_march_kernel<32, 58, NS2>;
    const size_t per_warp = NS2 * mpd_dev::march_warp_doubles(g.nx) * sizeof(double);
    const int cap = 12 / NS2;  // the kernel's __launch_bounds__: 12 warps of 162 registers / 6 warps of up to 255
    int warps = g.march_warps > 0 ? g.march_warps : default_warps > 0 ? default_warps : cap;
    if (warps > cap) warps = cap;
    if ((size_t)warps * per_warp > (size_t)g.smem_optin) warps = (int)((size_t)g.smem_optin / per_warp);
    if (warps < 1) return fail(3, "the rings of one warp need %zu B of shared memory (nx=%d), the SM has %d", per_warp, g.nx, g.smem_optin);
    const size_t smem = (size_t)warps * per_warp;
    CU(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, warps * 32, smem));
    if (per_sm < 1) return fail(3, "the marching form does not fit an SM with %d warps per CTA", warps);
    const long long per_cta = (long long)warps * NS2, ctas = ((long long)g.ns + per_cta - 1) / per_cta, resident = (long long)per_sm * g.sms;
    g.use_form = 1 + NS2; g.use_chunks = 1; g.use_block = warps * 32; g.nlmax = g.nz - 1; g.smem = smem; g.prefetch = 0;
    g.grid = (int)(ctas < resident ? ctas : resident);
    return 0;
}

int plan_launch() {
    const int nzm = g.nz - 1;
    {
        int form = g.form;
        if (form == 0)
            if (const char *env = getenv("MPDATA_B200_FORM")) form = atoi(env);
        const bool ref_sizes = g.nx == 32 && g.nz == 58;  // the sizes the marching form is instantiated for (A:6-8)
        if (form >= 2 && !ref_sizes) return fail(2, "the marching form is compiled for nx=32, nz=58 only (got nx=%d, nz=%d)", g.nx, g.nz);
        if (form == 2) return plan_march<1>();
        if (form == 3) return plan_march<2>();
        // Automatic choice (form 0, no launch-shape override) at the reference plane sizes: the marching kernel with 6 warps
        // per CTA where march_wins() expects it to be the faster one.  Both forms give the oracle's bits.
        if (form == 0 && ref_sizes && g.chunks <= 0 && g.block <= 0 && !g.force_generic) {
            int on = 1;
            if (const char *env = getenv("MPDATA_B200_AUTO_MARCH")) on = atoi(env);
            if (on && mpdata_march_wins(g.ns, g.sms)) return plan_march<1>(6);
        }
        g.use_form = 1;
    }
    int chunks = g.chunks;
    if (chunks <= 0) {
        chunks = g.sms / g.ns;
        if (chunks > nzm / 8) chunks = nzm / 8;
        // at least as many slices as SMs: two half-slice CTAs per SM overlap each other's barriers and load phases
        // (measured: 1,184 slices 0.160 against 0.171 ms, 9,472 slices 1.23 against 1.29 ms; profiles/r2a_mpdata_prof.json)
        if (chunks < 2 && g.ns >= g.sms && nzm >= 16) chunks = 2;
    }
    if (chunks > nzm) chunks = nzm;
    if (chunks < 1) chunks = 1;
    int nlmax = mpd_dev::max_local_levels(nzm, chunks);
    size_t smem = mpd_dev::item_doubles(g.nx, nlmax) * sizeof(double);
    while (smem > (size_t)g.smem_optin && chunks < nzm) {  // a whole slice does not fit: more, shorter chunks
        ++chunks;
        nlmax = mpd_dev::max_local_levels(nzm, chunks);
        smem = mpd_dev::item_doubles(g.nx, nlmax) * sizeof(double);
    }
    if (smem > (size_t)g.smem_optin)
        return fail(3, "one level range of a slice needs %zu B of shared memory (nx=%d), the SM has %d", smem, g.nx, g.smem_optin);
    int block = g.block;
    if (block <= 0) {
        const long long pts = (long long)(g.nx + 6) * nlmax;
        block = pts >= 1536 ? 1024 : pts >= 640 ? 512 : pts >= 192 ? 256 : 64;
    }
    g.use_chunks = chunks; g.use_block = block; g.nlmax = nlmax; g.smem = smem;
    // the kernel is compiled for 64 registers per thread: 64 K registers hold 1024 threads per SM
    int per_sm = (int)((size_t)g.smem_sm / (smem + 1024));
    if (per_sm > 1024 / block) per_sm = 1024 / block;
    if (per_sm < 1) per_sm = 1;
    const long long items = (long long)g.ns * chunks, resident = (long long)per_sm * g.sms;
    g.grid = (int)(items < resident ? items : resident);
    // Prefetching form (MPDATA_B200_PREFETCH=1, off by default): when a CTA walks several items and a second set of the input
    // planes fits without costing a resident CTA, the next item's f, u, w and level columns travel (cp.async) while the
    // current item is computed.  Measured (profiles/r3h_*): bit-identical, but no gain at the default shape (9,472 slices: 1.27
    // against 1.22 ms with two half-slice CTAs per SM; it only helps one CTA per SM, 1.24 against 1.30 ms) -- the global-load
    // phase is not what bounds this kernel.
    g.prefetch = 0;
    {
        const size_t smem2 = smem + mpd_dev::prefetch_doubles(g.nx, nlmax) * sizeof(double);
        int per_sm2 = (int)((size_t)g.smem_sm / (smem2 + 1024));
        if (per_sm2 > 1024 / block) per_sm2 = 1024 / block;
        const char *env = getenv("MPDATA_B200_PREFETCH");
        const bool want = env ? atoi(env) != 0 : false;
        if (want && items > g.grid && smem2 <= (size_t)g.smem_optin && per_sm2 >= per_sm) {
            g.prefetch = 1;
            smem = smem2;
            g.smem = smem2;
        }
    }
    CU(cudaFuncSetAttribute(mpd_dev::advect_kernel<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CU(cudaFuncSetAttribute(mpd_dev::advect_kernel<32, 58>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return 0;
}

int launch_step() {
    // f -> f1, then the two swap roles: a level chunk's halo must read the other chunks' input
    mpd_dev::Problem p{g.ns, g.nx, g.nz, g.f, g.f1, g.u, g.w, g.rho, g.rhow, g.adz, g.flux};
    if (g.use_form == 2)
        mpd_dev::advect_march_kernel<32, 58, 1><<<g.grid, g.use_block, g.smem, g.stream>>>(p);
    else if (g.use_form == 3)
        mpd_dev::advect_march_kernel<32, 58, 2><<<g.grid, g.use_block, g.smem, g.stream>>>(p);
    else if (g.nx == 32 && g.nz == 58 && !g.force_generic)  // the reference sizes (A:6-8): plane sizes as compile-time constants
        mpd_dev::advect_kernel<32, 58><<<g.grid, g.use_block, g.smem, g.stream>>>(p, g.use_chunks, g.nlmax, g.prefetch);
    else
        mpd_dev::advect_kernel<0, 0><<<g.grid, g.use_block, g.smem, g.stream>>>(p, g.use_chunks, g.nlmax, g.prefetch);
    ++g_launches;
    CU(cudaGetLastError());
    double *t = g.f; g.f = g.f1; g.f1 = t;
    return 0;
}

}  // namespace

extern "C" {

const char *mpdata_last_error(void) { return g_failed ? g_err : nullptr; }
long long mpdata_launch_count(void) { return g_launches; }
long long mpdata_slice_smem_bytes(int nx, int nz) {
    if (nx < 1 || nz < 3) return -1;
    return (long long)(mpd_dev::item_doubles(nx, nz - 1) * sizeof(double));
}

int mpdata_set_block(int threads) {
    if (threads != 0 && (threads < 32 || threads > 1024 || threads % 32))
        return fail(2, "mpdata_set_block: %d is not 0 (automatic) or a multiple of 32 in 32..1024", threads);
    g.block = threads;
    return g.ready ? plan_launch() : 0;
}

int mpdata_set_chunks(int chunks) {
    if (chunks < 0) return fail(2, "mpdata_set_chunks: %d is negative (0 = automatic)", chunks);
    g.chunks = chunks;
    return g.ready ? plan_launch() : 0;
}

int mpdata_set_form(int form, int warps_per_cta) {
    if (form < 0 || form > 3 || warps_per_cta < 0 || warps_per_cta > 12)
        return fail(2, "mpdata_set_form: form %d is not 0 (automatic), 1 (phases), 2 or 3 (marching, one / two slices per warp), or %d warps per CTA is not in 0..12", form, warps_per_cta);
    const int form0 = g.form, warps0 = g.march_warps;
    g.form = form;
    g.march_warps = warps_per_cta;
    if (!g.ready) return 0;
    const int rc = plan_launch();
    if (rc) {  // e.g. the marching form asked for while a context of other plane sizes is live: nothing changes
        g.form = form0;
        g.march_warps = warps0;
        plan_launch();
    }
    return rc;
}

int mpdata_get_launch(int *grid, int *block, int *chunks, long long *smem_bytes) {
    if (!g.ready) return fail(5, "mpdata_get_launch before mpdata_init");
    if (grid) *grid = g.grid;
    if (block) *block = g.use_block;
    if (chunks) *chunks = g.use_chunks;
    if (smem_bytes) *smem_bytes = (long long)g.smem;
    return 0;
}

int mpdata_init(int nslices, int nx, int nz) {
    g_failed = false;
    if (nslices < 1 || nx < 1 || nz < 3) return fail(2, "mpdata_init: bad sizes nslices=%d nx=%d nz=%d (need >=1, >=1, >=3)", nslices, nx, nz);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
        return fail(1, "mpdata_init: no CUDA device (this library has no CPU fallback)");
    int dev = 0;
    if (const char *lr = getenv("LOCAL_RANK")) dev = atoi(lr) % ndev;
    CU(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10) return fail(1, "mpdata_init: device %d is sm_%d%d, this library is built for sm_100a only", dev, prop.major, prop.minor);
    if (mpd_dev::item_doubles(nx, nz - 1 < 2 * mpd_dev::kHalo + 1 ? nz - 1 : 2 * mpd_dev::kHalo + 1) * sizeof(double) >
        (size_t)prop.sharedMemPerBlockOptin)
        return fail(3, "mpdata_init: even one level of a slice with its halo does not fit shared memory (nx=%d): %zu B, the SM has %zu",
                    nx, mpd_dev::item_doubles(nx, 2 * mpd_dev::kHalo + 1) * sizeof(double), (size_t)prop.sharedMemPerBlockOptin);
    release();
    g.force_generic = getenv("MPDATA_B200_GENERIC") && atoi(getenv("MPDATA_B200_GENERIC")) != 0;
    g.ns = nslices; g.nx = nx; g.nz = nz;
    const size_t nzm = nz - 1;
    g.nf = (size_t)nslices * (nx + 6) * nzm;
    g.nu = (size_t)nslices * (nx + 5) * nzm;
    g.nw = (size_t)nslices * (nx + 4) * nz;
    g.n2m = (size_t)nslices * nzm;
    g.n2 = (size_t)nslices * nz;
    g.sms = prop.multiProcessorCount;
    g.smem_optin = (int)prop.sharedMemPerBlockOptin;
    g.smem_sm = (int)prop.sharedMemPerMultiprocessor;
    if (int rc = plan_launch()) return rc;
    CU(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
    CU(cudaMalloc(&g.f, g.nf * 8));
    CU(cudaMalloc(&g.f1, g.nf * 8));
    CU(cudaMalloc(&g.f0, g.nf * 8));
    CU(cudaMalloc(&g.u, g.nu * 8));
    CU(cudaMalloc(&g.w, g.nw * 8));
    CU(cudaMalloc(&g.rho, g.n2m * 8));
    CU(cudaMalloc(&g.rhow, g.n2 * 8));
    CU(cudaMalloc(&g.adz, g.n2m * 8));
    CU(cudaMalloc(&g.flux, g.n2 * 8));
    CU(cudaMemsetAsync(g.flux, 0, g.n2 * 8, g.stream));
    g.ready = true;
    return 0;
}

int mpdata_upload(const double *f, const double *u, const double *w, const double *rho, const double *rhow,
                  const double *adz, const double *flux) {
    if (!g.ready) return fail(5, "mpdata_upload before mpdata_init");
    if (!f || !u || !w || !rho || !rhow || !adz) return fail(2, "mpdata_upload: null array");
    CU(cudaMemcpyAsync(g.f, f, g.nf * 8, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.u, u, g.nu * 8, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.w, w, g.nw * 8, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.rho, rho, g.n2m * 8, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.rhow, rhow, g.n2 * 8, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.adz, adz, g.n2m * 8, cudaMemcpyHostToDevice, g.stream));
    if (flux) CU(cudaMemcpyAsync(g.flux, flux, g.n2 * 8, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.f0, g.f, g.nf * 8, cudaMemcpyDeviceToDevice, g.stream));
    g.uploaded = true;
    return 0;
}

int mpdata_run(void) {
    if (!g.ready || !g.uploaded) return fail(5, "mpdata_run before mpdata_init/mpdata_upload");
    return launch_step();
}

int mpdata_download(double *f, double *flux) {
    if (!g.ready || !g.uploaded) return fail(5, "mpdata_download before mpdata_init/mpdata_upload");
    if (f) CU(cudaMemcpyAsync(f, g.f, g.nf * 8, cudaMemcpyDeviceToHost, g.stream));
    if (flux) CU(cudaMemcpyAsync(flux, g.flux, g.n2m * 8, cudaMemcpyDeviceToHost, g.stream));  // levels 1..nzm only
    CU(cudaStreamSynchronize(g.stream));
    return 0;
}

int mpdata_finalize(void) { return release(); }

int mpdata_advect_scalar2D(int nslices, int nx, int nz, double *f, const double *u, const double *w, const double *rho,
                           const double *rhow, const double *adz, double *flux) {
    if (!f || !flux) return fail(2, "mpdata_advect_scalar2D: null array");
    if (!g.ready || g.ns != nslices || g.nx != nx || g.nz != nz) {
        if (int rc = mpdata_init(nslices, nx, nz)) return rc;
    }
    if (int rc = mpdata_upload(f, u, w, rho, rhow, adz, nullptr)) return rc;
    if (int rc = launch_step()) return rc;
    return mpdata_download(f, flux);
}

int mpdata_time_steps(int iters, int flush_l2, float *ms_per_step) {
    if (!g.ready || !g.uploaded) return fail(5, "mpdata_time_steps before mpdata_init/mpdata_upload");
    if (iters < 1 || !ms_per_step) return fail(2, "mpdata_time_steps: bad arguments");
    if (flush_l2 && !g.flush) {
        g.flush_bytes = (size_t)256 << 20;  // > the 126 MB L2
        CU(cudaMalloc(&g.flush, g.flush_bytes));
    }
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a));
    CU(cudaEventCreate(&b));
    double total = 0.0;
    for (int it = 0; it < iters; ++it) {
        CU(cudaMemcpyAsync(g.f, g.f0, g.nf * 8, cudaMemcpyDeviceToDevice, g.stream));
        if (flush_l2) CU(cudaMemsetAsync(g.flush, it & 0xff, g.flush_bytes, g.stream));
        CU(cudaEventRecord(a, g.stream));
        if (int rc = launch_step()) return rc;
        CU(cudaEventRecord(b, g.stream));
        CU(cudaEventSynchronize(b));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, a, b));
        total += ms;
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    *ms_per_step = (float)(total / iters);
    return 0;
}

}  // extern "C"
