// bwk_api.cu -- the C ABI of include/bwk_b200.h on top of the sm_100a kernels.
//
// Mirrors the reference driver's data life cycle (atmosphere/biharmonic_wk_kernel.F90):
//   allocate_data + enter data pcreate  F90:37-44,553  -> bwk_init
//   update device                        F90:573        -> bwk_upload
//   bws_gpu2(...)                        F90:574        -> bwk_run
//   update host                          F90:575        -> bwk_download
// There is no CPU fallback: every compute entry point needs a CUDA device.
#include "../../include/bwk_b200.h"

#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "bwk_internal.h"
#include "bwk_kernels.h"

using namespace bwk;

namespace {

thread_local std::string g_err;

int fail(const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

#define CK(call)                                                                                            \
    do {                                                                                                    \
        cudaError_t e_ = (call);                                                                            \
        if (e_ != cudaSuccess) return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ---- NCCL through dlopen: single-GPU users need no NCCL at all
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool load()
    {
        if (lib) return true;
        lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return false;
#define SYM(name) *(void **)(&name) = dlsym(lib, "nccl" #name)
        SYM(GetUniqueId);
        SYM(CommInitRank);
        SYM(CommDestroy);
        SYM(Send);
        SYM(Recv);
        SYM(GroupStart);
        SYM(GroupEnd);
        SYM(GetErrorString);
#undef SYM
        return GetUniqueId && CommInitRank && CommDestroy && Send && Recv && GroupStart && GroupEnd && GetErrorString;
    }
} g_nccl;

#define NK(call)                                                                                              \
    do {                                                                                                      \
        ncclResult_t r_ = (call);                                                                             \
        if (r_ != ncclSuccess) return fail("%s failed: %s (%s:%d)", #call, g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
    } while (0)

constexpr int NPARTIAL_BLOCKS = 148 * 8;
constexpr int NEVENTS = 64;

// the warp-strip sweep's views of a field laid out like qtens: boxes of 32/NM slabs x NM elements, NM = 4 << i
struct FieldMaps {
    CUtensorMap slab[3], rows[3];
};
constexpr int WS_NM[3] = {4, 8, 16};

struct Ctx {
    bool inited = false;
    int nelem = 0, nlev = 0, qsize = 0, kq = 0;
    size_t q_doubles = 0;
    cudaStream_t s_main = nullptr, s_comm = nullptr, s_dep = nullptr, s_aux = nullptr, s_h2d = nullptr, s_d2h = nullptr;
    double *d_q = nullptr, *d_tmp = nullptr, *d_elem = nullptr, *d_metric = nullptr, *d_rsph = nullptr;
    double *d_w = nullptr, *d_ratio = nullptr; // advance_hypervis_scalar: biharmonic work buffer, dpdiss/dp [nelem][nlev][16]
    CUtensorMap map_w, map_w_rows, map_w_pt; // the work buffer seen as slabs / edge rows / i-pairs (strip sweeps reading it)
    FieldMaps fm_tmp, fm_w, fm_q;            // warp-strip sweep: views of tmp, the work buffer and qtens
    CUtensorMap map_halo_nm[3];              // ... and of the received rows
    StripReg *d_grp[3][3] = {};              // [list][i]: groups of WS_NM[i]/4 consecutive fast strips, records back to back
    int64_t ngrp[3][3] = {};
    bool src_is_w = false;                  // strip sweeps and the halo pack read the work buffer instead of tmp
    double *d_v2c = nullptr, *d_metric_plain = nullptr; // biharmonic_wk_dp3d: vec_sphere2cart [nelem][96]; folded metric without tensorVisc
    bool have_v2c = false, have_metric_plain = false;
    double *d_save = nullptr, *d_table = nullptr, *d_partials = nullptr, *d_halo = nullptr, *d_send = nullptr, *d_dummy = nullptr;
    double *h_partials = nullptr;
    ElemDss *d_dss = nullptr;
    int32_t *d_boundary = nullptr, *d_interior = nullptr, *d_dependent = nullptr, *d_independent = nullptr;
    int32_t *d_send_elem = nullptr, *d_send_row = nullptr;
    int32_t *d_strips_all = nullptr, *d_strips_dep = nullptr, *d_strips_indep = nullptr;
    StripReg *d_reg_all = nullptr, *d_reg_dep = nullptr, *d_reg_indep = nullptr;
    int *d_work = nullptr;            // item counter of the persistent strip sweep
    int64_t nfast[3] = {0, 0, 0};     // leading strips of the all / dependent / independent lists the persistent sweep takes (full, regular)
    std::vector<int64_t> send_offset; // per peer, rows
    bool have_metric = false, have_topology = false, have_rsph = false, rsph_user = false;
    SweepParams sp{};
    Plan plan;
    std::vector<double> h_spheremp; // spheremp of the local elements (16 each), for rspheremp
    int64_t gen_nelem_global = -1;  // >=0 when the data came from bwk_initialize_data_device (LCG closed form)
    int gen_seed_mode = 0;
    int64_t gen_elem_first = -1;
    int variant = 0;
    int phase_event = -1;             // bwk_set_phase_event: event slot recorded between sweep 1 and the strip sweep (1 rank)
    int64_t launches = 0;
    // comm
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1;
    cudaEvent_t ev_b = nullptr, ev_halo = nullptr, ev_misc = nullptr, ev_start = nullptr, ev_int = nullptr, ev_aux0 = nullptr, ev_aux1 = nullptr;
    cudaEvent_t events[NEVENTS] = {};
    std::vector<cudaEvent_t> chunk_up, chunk_done;
    // host-buffer pipeline of the full biharmonic (bwk_biharmonic_wk_scalar_full): strips_all re-ordered
    // by the upload chunk that completes their inputs, see build_e2e_schedule()
    bool e2e_ready = false;
    std::vector<int> e2e_bounds;                           // upload chunk boundaries (elements)
    std::vector<int> e2e_group_start;                      // [nchunks+1] first strip (in the re-ordered list) of each group
    std::vector<std::vector<std::pair<int, int>>> e2e_runs; // per group: contiguous element runs [first, end) to download
    int32_t *d_e2e_strips = nullptr;
    StripReg *d_e2e_reg = nullptr;
};
Ctx g;

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

int get_encode()
{
    if (g_encode) return 0;
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (qres != cudaDriverEntryPointSuccess || !fn) return fail("cuTensorMapEncodeTiled is not available from this driver");
    g_encode = (EncodeTiledFn)fn;
    return 0;
}

int make_map(CUtensorMap *m, void *base, int rank, const uint64_t *dims, const uint64_t *strides_bytes /*rank-1*/, const uint32_t *box,
             CUtensorMapSwizzle swz, CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_L2_128B)
{
    if (get_encode()) return 1;
    cuuint64_t gd[5], gs[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
    CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, (cuuint32_t)rank, base, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                          promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu %llu %llu)", (int)r, rank,
                                       (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)(rank > 2 ? dims[2] : 0));
    return 0;
}

// {16, KQ, nelem} slabs, box {16,32,1}, SWIZZLE_128B
int make_slab_map(CUtensorMap *m, double *base, int kq, int64_t nelem)
{
    const uint64_t dims[3] = {16, (uint64_t)kq, (uint64_t)std::max<int64_t>(nelem, 1)};
    const uint64_t str[2] = {128, (uint64_t)kq * 128};
    const uint32_t box[3] = {16, TILE_SLABS, 1};
    return make_map(m, base, 3, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_field_maps(FieldMaps *m, double *base, int kq, int64_t nelem)
{
    const uint64_t n = (uint64_t)std::max<int64_t>(nelem, 1);
    const uint64_t dims[3] = {16, (uint64_t)kq, n}, str[2] = {128, (uint64_t)kq * 128};
    for (int i = 0; i < 3; ++i) {
        const uint32_t nm = (uint32_t)std::min<uint64_t>((uint64_t)WS_NM[i], n), sl = 32u / (uint32_t)WS_NM[i];
        const uint32_t b_slab[3] = {16, sl, nm}, b_rows[3] = {4, sl, nm};
        if (make_map(&m->slab[i], base, 3, dims, str, b_slab, CU_TENSOR_MAP_SWIZZLE_128B)) return 1;
        if (make_map(&m->rows[i], base, 3, dims, str, b_rows, CU_TENSOR_MAP_SWIZZLE_32B)) return 1;
    }
    return 0;
}

int need_init()
{
    if (!g.inited) return fail("bwk_init has not been called");
    return 0;
}

int check_range(int nets, int nete)
{
    if (nets < 1 || nete > g.nelem || nets > nete + 1) return fail("element range nets=%d nete=%d outside 1..%d", nets, nete, g.nelem);
    return 0;
}

// launch one element-local sweep (MODE_LAPLACE / MODE_SWEEP1) over local elements
// [first, first+count) or over a device list
int sweep(int mode, int first, int count, const int32_t *list, cudaStream_t st)
{
    if (count <= 0) return 0;
    SweepParams p = g.sp;
    const SweepGeometry geo = choose_geometry(g.kq, count, mode);
    p.tiles_per_unit = geo.tiles_per_unit;
    p.units = geo.units;
    p.elem_list = list;
    p.elem_first = first;
    p.elem_count = count;
    CK(launch_sweep(mode, p, st));
    g.launches++;
    return 0;
}

// Unit length of the strip sweeps, tuned per launch shape.  The best number of tiles per CTA depends on how long the
// neighbouring CTAs stay in step (L2 residency of shared lines) against the per-CTA pipeline fill, and that moves with the
// strip count and the mesh (measured: 15 tiles best at ne120/ne60/ne30 on one rank, 9 tiles 6-8 % faster at ne84, ne100 and
// on the per-rank blocks of a 2- or 8-rank run).  So the first calls of a shape measure both with CUDA events (results do
// not depend on the unit length): call 0 runs candidate 0 untimed (cold), calls 1 and 2 time candidates 0 and 1, later
// calls use the faster one.
struct StripTune {
    int calls = 0, best = -1;
    cudaEvent_t ev[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
};
constexpr int TUNE_CAND[2] = {15, 9};
// warp-strip sweep with NM members per warp: a CTA walks NM/4 times as many (smaller) steps per tile and lives as much longer,
// and DRAM reads grow with the CTA lifetime (neighbouring CTAs drift apart): 8 members at ne120: 5 tiles 11.3 ms / 1.06x reads,
// 7: 11.2 ms / 1.13x, 9: 11.4 ms / 1.19x, 15: 12.9 ms / 1.43x (profiles/r2y_*)
inline int tune_cand(int members, int c) { return members >= 16 ? (c ? 3 : 4) : members >= 8 ? (c ? 5 : 7) : TUNE_CAND[c]; }
std::map<uint64_t, StripTune> g_tune;

// launch sweep 2 (or the DSS-only sweep) over `count` strips of a device strip list
int sweep2(const int32_t *d_strips, const StripReg *d_reg, int64_t count, cudaStream_t st, int mode = MODE_SWEEP2, const CUtensorMap *out = nullptr,
           int members = 0 /* > 0: warp-strip sweep over `count` groups of members/4 strips */)
{
    if (count <= 0) return 0;
    SweepParams p = g.sp;
    if (out) p.map_out = *out;
    if (g.src_is_w) {
        p.tmp = g.d_w;
        p.map_tmp = g.map_w;
        p.map_rows = g.map_w_rows;
        p.map_pt = g.map_w_pt;
    }
    if (members) { // the same choices for the warp-strip sweep's views (boxes of 32/members slabs x members elements)
        static const int dbg = std::getenv("BWK_DEBUG_FLAGS") ? std::atoi(std::getenv("BWK_DEBUG_FLAGS")) : 0;
        p.debug_flags = dbg;
        const int i = members == 16 ? 2 : members == 8 ? 1 : 0;
        const FieldMaps &src = g.src_is_w ? g.fm_w : g.fm_tmp;
        p.map4_tmp = src.slab[i];
        p.map4_rows = src.rows[i];
        p.map4_halo = g.map_halo_nm[i];
        p.map4_out = out == &g.map_w ? g.fm_w.slab[i] : g.fm_q.slab[i];
        p.map4_q = g.fm_q.slab[i];
    }
    int hint = 0, timed = -1;
    StripTune *t = nullptr;
    if (sweep2_tiles_per_unit_forced() == 0 && count * ((g.kq + TILE_SLABS - 1) / TILE_SLABS) >= 148 * 64 && g_tune.size() < 4096) {
        t = &g_tune[((uint64_t)mode << 56) ^ ((uint64_t)members << 50) ^ ((uint64_t)count << 20) ^ (uint64_t)g.kq];
        if (t->best >= 0) hint = tune_cand(members, t->best);
        else if (t->calls <= 1) { hint = tune_cand(members, 0); timed = t->calls == 1 ? 0 : -1; }
        else if (t->calls == 2) { hint = tune_cand(members, 1); timed = 1; }
        else {
            float ms[2] = {0.f, 0.f};
            for (int c = 0; c < 2; ++c) {
                CK(cudaEventSynchronize(t->ev[c][1]));
                CK(cudaEventElapsedTime(&ms[c], t->ev[c][0], t->ev[c][1]));
                CK(cudaEventDestroy(t->ev[c][0]));
                CK(cudaEventDestroy(t->ev[c][1]));
                t->ev[c][0] = t->ev[c][1] = nullptr;
            }
            t->best = ms[1] < 0.975f * ms[0] ? 1 : 0; // single samples: the default keeps the shape unless the alternative wins clearly
            hint = tune_cand(members, t->best);
        }
        t->calls++;
    }
    if (hint == 0 && members) hint = tune_cand(members, 0);
    const SweepGeometry geo = choose_geometry(g.kq, count, mode, hint);
    p.tiles_per_unit = geo.tiles_per_unit;
    p.units = geo.units;
    p.strips = d_strips;
    p.strip_reg = d_reg;
    p.strip_count = (int)count;
    if (timed >= 0) {
        CK(cudaEventCreate(&t->ev[timed][0]));
        CK(cudaEventCreate(&t->ev[timed][1]));
        CK(cudaEventRecord(t->ev[timed][0], st));
    }
    CK(members ? launch_sweep2_wstrip(mode, p, st, members) : launch_sweep(mode, p, st));
    if (timed >= 0) CK(cudaEventRecord(t->ev[timed][1], st));
    g.launches++;
    return 0;
}

// strip sweep over one of the plan's three lists (0 all, 1 dependent, 2 independent): the full regular strips through the
// persistent kernel (units of BWK_PERSIST_T tiles, default 4), the others through sweep2_strip_kernel behind it
int sweep2_list(int which, cudaStream_t st, int mode = MODE_SWEEP2, const CUtensorMap *out = nullptr)
{
    int32_t *const ds[3] = {g.d_strips_all, g.d_strips_dep, g.d_strips_indep};
    StripReg *const dr[3] = {g.d_reg_all, g.d_reg_dep, g.d_reg_indep};
    const size_t sizes[3] = {g.plan.strips_all.size(), g.plan.strips_dependent.size(), g.plan.strips_independent.size()};
    const int64_t n = (int64_t)sizes[which];
    int64_t nf = g.nfast[which];
    // A/B knob BWK_STRIP_FORM: 2 (default) = warp-strip sweep for the fast strips, 1 = persistent one-warp-per-member sweep,
    // 0 = everything through sweep2_strip_kernel (the round-1 form)
    static const int form = std::getenv("BWK_STRIP_FORM") ? std::atoi(std::getenv("BWK_STRIP_FORM")) : 2;
    static const int tpu = std::getenv("BWK_PERSIST_T") ? std::max(3, std::atoi(std::getenv("BWK_PERSIST_T"))) : 4;
    const int ntiles = (g.kq + TILE_SLABS - 1) / TILE_SLABS;
    // groups of chained strips pay on long lists only (measured: -4..6 % on the full ne120 sphere, +2 % on a rank's block of an
    // 8-rank run, where the halved CTA count costs more than the halved strip ends give): below BWK_WSTRIP_GROUP_MIN fast
    // strips every strip runs alone (4 members per warp)
    static const int64_t group_min = std::getenv("BWK_WSTRIP_GROUP_MIN") ? std::atoll(std::getenv("BWK_WSTRIP_GROUP_MIN")) : 4096;
    if (form == 2 && nf > 0) {
        // The strips the warp-strip sweep does not take (1.1 % of them on the ne120 sphere: face-edge strips with reversed rows,
        // short strips) go through sweep2_strip_kernel on a side stream, launched FIRST: their few CTAs run at full occupancy
        // while the main grid fills in behind them, instead of as a 3-wave tail of their own (0.43 ms = 3.8 % of the sweep).
        const bool side = n > nf;
        if (side) {
            CK(cudaEventRecord(g.ev_aux0, st));
            CK(cudaStreamWaitEvent(g.s_aux, g.ev_aux0, 0));
            if (sweep2(ds[which] + nf, dr[which] + nf, n - nf, g.s_aux, mode, out)) return 1;
            CK(cudaEventRecord(g.ev_aux1, g.s_aux));
        }
        if (nf < group_min) {
            if (sweep2(nullptr, dr[which], nf, st, mode, out, 4)) return 1;
        } else {
            for (int gi = 2; gi >= 0; --gi) {
                if (g.ngrp[which][gi] == 0) continue;
                // (SK_POST keeps two more arrays per member in shared memory: 16 members do not fit)
                if (mode == MODE_SWEEP2_POST && WS_NM[gi] == 16) return fail("advance_hypervis_scalar with BWK_WSTRIP_MEMBERS=16 is not supported");
                if (sweep2(nullptr, g.d_grp[which][gi], g.ngrp[which][gi], st, mode, out, WS_NM[gi])) return 1;
            }
        }
        if (side) CK(cudaStreamWaitEvent(st, g.ev_aux1, 0));
        return 0;
    }
    if (form != 1 || ntiles < 3 || (mode != MODE_SWEEP2 && mode != MODE_DSS)) nf = 0;
    if (nf > 0) {
        SweepParams p = g.sp;
        if (out) p.map_out = *out;
        if (g.src_is_w) {
            p.tmp = g.d_w;
            p.map_tmp = g.map_w;
            p.map_rows = g.map_w_rows;
            p.map_pt = g.map_w_pt;
        }
        p.strip_reg = dr[which];
        p.strip_count = (int)nf;
        p.work_counter = g.d_work;
        static const int nbr_ahead = std::getenv("BWK_NBR_AHEAD") ? std::atoi(std::getenv("BWK_NBR_AHEAD")) : 2; // tuning knob
        p.nbr_ahead = nbr_ahead == 1 ? 1 : 2;
        CK(cudaMemsetAsync(g.d_work, 0, sizeof(int), st));
        p.units = std::max(1, ntiles / tpu);
        p.tiles_per_unit = ntiles / p.units;
        CK(launch_sweep2_persist(mode, p, st));
        g.launches++;
    }
    return sweep2(ds[which] + nf, dr[which] + nf, n - nf, st, mode, out);
}

int laplace_range(int first0, int count, cudaStream_t st)
{
    if (!g.have_metric) return fail("no element metric on the device: call bwk_upload(elem, ..., deriv) first");
    if (g.variant == 1) {
        CK(launch_laplace_plain(g.d_q, g.d_metric, first0, count, g.kq, st));
        g.launches += (count + 65534) / 65535;
        return 0;
    }
    return sweep(MODE_LAPLACE, first0, count, nullptr, st);
}

int upload_metric(const bwk_element_t *elem, const bwk_derivative_t *deriv, cudaStream_t st)
{
    if (elem) {
        CK(cudaMemcpyAsync(g.d_elem, elem, sizeof(double) * ELEM_DOUBLES * g.nelem, cudaMemcpyHostToDevice, st));
        CK(launch_fold_metric(g.d_elem, g.d_metric, g.nelem, st));
        g.launches++;
        g.have_metric = true;
        g.have_metric_plain = false;
        g.gen_nelem_global = -1;
        const double *e = reinterpret_cast<const double *>(elem);
        bool same = g.h_spheremp.size() == (size_t)16 * g.nelem;
        g.h_spheremp.resize((size_t)16 * g.nelem);
        for (int64_t ie = 0; ie < g.nelem; ++ie) {
            double *dst = &g.h_spheremp[(size_t)16 * ie];
            const double *src = e + ie * ELEM_DOUBLES + 64;
            if (same && std::memcmp(dst, src, 16 * sizeof(double)) != 0) same = false;
            std::memcpy(dst, src, 16 * sizeof(double));
        }
        // rspheremp = 1/DSS(spheremp) stays valid while spheremp is unchanged (HOMME computes it once at
        // start-up); a caller-supplied rspheremp (bwk_set_rspheremp) stays in force regardless
        if (!g.rsph_user && !same) g.have_rsph = false;
    }
    if (deriv) CK(set_dvv(&deriv->Dvv[0][0], st));
    return 0;
}

void free_dev(void *p)
{
    if (p) cudaFree(p);
}

std::vector<double> lcg_table()
{
    std::vector<double> t(131072);
    int old = 11; // F90:83-89
    for (int i = 0; i < 131072; ++i) {
        old = (1301 * old + 97) % (1024 * 128);
        t[(size_t)i] = old / (double)(1024 * 128);
    }
    return t;
}

// rspheremp = 1 / DSS(spheremp); needs spheremp of remote sharers when nranks > 1
int compute_rspheremp()
{
    if (g.have_rsph) return 0;
    if (!g.have_topology) return fail("no topology set");
    if ((int64_t)g.h_spheremp.size() != (int64_t)16 * g.nelem) return fail("no element metric uploaded/generated before the topology was used");
    const Plan &pl = g.plan;
    if (pl.nranks > 1 && g.gen_nelem_global < 0) {
        // Caller-supplied element records on several ranks (a real mesh): the sharers' spheremp on other ranks is not known
        // here, so DSS it through the halo plan -- the spheremp values of exactly the rows the plan sends travel once with
        // ncclSend/Recv (4 doubles per row), and the plan's own gather, evaluated on the host, gives DSS(spheremp) in the
        // kernels' summation order.  Collective: every rank of the communicator must get here (same data source everywhere).
        if (!g.comm) return fail("topology spans %d ranks but bwk_comm_init was not called", pl.nranks);
        CK(cudaDeviceSynchronize());
        const int64_t ns = g.send_offset.empty() ? 0 : g.send_offset.back(), nr = pl.recv_offset.back();
        std::vector<double> hs((size_t)4 * std::max<int64_t>(ns, 1)), hr((size_t)4 * std::max<int64_t>(nr, 1));
        auto row_points = [](int row, int t) { return row < 4 ? t + 4 * row : (row == ROW_W ? 0 : 3) + 4 * t; };
        {
            size_t k = 0;
            for (int p = 0; p < pl.nranks; ++p)
                for (const RowRef &r : pl.send[(size_t)p]) {
                    const double *sp = &g.h_spheremp[(size_t)16 * (r.elem - pl.first)];
                    for (int t = 0; t < 4; ++t) hs[4 * k + t] = sp[row_points(r.row, t)];
                    ++k;
                }
        }
        if (ns > 0) CK(cudaMemcpyAsync(g.d_send, hs.data(), sizeof(double) * 4 * ns, cudaMemcpyHostToDevice, g.s_comm));
        NK(g_nccl.GroupStart());
        for (int p = 0; p < pl.nranks; ++p) {
            if (p == pl.rank) continue;
            const int64_t a = g.send_offset[(size_t)p + 1] - g.send_offset[(size_t)p], b = pl.recv_offset[(size_t)p + 1] - pl.recv_offset[(size_t)p];
            if (a > 0) NK(g_nccl.Send(g.d_send + 4 * g.send_offset[(size_t)p], (size_t)4 * a, ncclDouble, p, g.comm, g.s_comm));
            if (b > 0) NK(g_nccl.Recv(g.d_halo + 4 * pl.recv_offset[(size_t)p], (size_t)4 * b, ncclDouble, p, g.comm, g.s_comm));
        }
        NK(g_nccl.GroupEnd());
        if (nr > 0) CK(cudaMemcpyAsync(hr.data(), g.d_halo, sizeof(double) * 4 * nr, cudaMemcpyDeviceToHost, g.s_comm));
        CK(cudaStreamSynchronize(g.s_comm));
        std::vector<double> sp_g((size_t)16 * pl.nelem_global, std::nan(""));
        std::memcpy(&sp_g[(size_t)16 * pl.first], g.h_spheremp.data(), sizeof(double) * 16 * g.nelem);
        for (int p = 0; p < pl.nranks; ++p)
            for (size_t i = 0; i < pl.recv[(size_t)p].size(); ++i) {
                const RowRef &r = pl.recv[(size_t)p][i];
                const double *v = &hr[(size_t)4 * (pl.recv_offset[(size_t)p] + (int64_t)i)];
                for (int t = 0; t < 4; ++t) sp_g[(size_t)16 * r.elem + row_points(r.row, t)] = v[t];
            }
        std::vector<double> rsx((size_t)16 * g.nelem);
        for (int64_t le = 0; le < g.nelem; ++le) {
            double sum16[16];
            gather_elem_host(pl, le, sp_g.data(), sum16);
            for (int t = 0; t < 16; ++t) {
                if (!(sum16[t] == sum16[t])) return fail("rspheremp: the halo plan did not deliver a sharer's spheremp (element %lld point %d)", (long long)le, t);
                rsx[(size_t)16 * le + t] = 1.0 / sum16[t];
            }
        }
        CK(cudaMemcpyAsync(g.d_rsph, rsx.data(), sizeof(double) * rsx.size(), cudaMemcpyHostToDevice, g.s_main));
        CK(cudaStreamSynchronize(g.s_main));
        g.have_rsph = true;
        return 0;
    }
    std::vector<double> table;
    if (g.gen_nelem_global >= 0 && g.gen_seed_mode == 0) table = lcg_table();
    auto remote_sp = [&](int64_t gp, double *out) -> bool { // generator data: a sharer's spheremp by closed form
        if (g.gen_nelem_global < 0) return false;
        const int64_t d = 16 + (int64_t)ELEM_DOUBLES * (gp / 16) + 64 + (gp % 16);
        if (g.gen_seed_mode == 0) { *out = table[(size_t)(d & 131071)]; return true; }
        uint64_t x = (uint64_t)d + 0x9E3779B97F4A7C15ull;
        x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
        x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
        x ^= x >> 31;
        *out = (double)(x >> 11) * 0x1.0p-53;
        return true;
    };
    std::vector<double> rs((size_t)16 * g.nelem);
    for (int64_t lp = 0; lp < (int64_t)16 * g.nelem; ++lp) {
        double s = 0.0;
        bool first = true;
        for (int64_t t = pl.pt_start[(size_t)lp]; t < pl.pt_start[(size_t)lp + 1]; ++t) {
            const int64_t gp = pl.pt_member[(size_t)t];
            const int64_t le = gp / 16 - pl.first;
            double v;
            if (le >= 0 && le < g.nelem) v = g.h_spheremp[(size_t)(16 * le + gp % 16)];
            else if (!remote_sp(gp, &v))
                return fail("rspheremp needs spheremp of elements owned by other ranks: call bwk_set_rspheremp (or generate the data with "
                            "bwk_initialize_data_device)");
            s = first ? v : s + v;
            first = false;
        }
        rs[(size_t)lp] = 1.0 / s;
    }
    CK(cudaMemcpyAsync(g.d_rsph, rs.data(), sizeof(double) * rs.size(), cudaMemcpyHostToDevice, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    g.have_rsph = true;
    return 0;
}

int upload_list(const std::vector<int32_t> &v, int32_t **d)
{
    free_dev(*d);
    *d = nullptr;
    if (v.empty()) return 0;
    CK(cudaMalloc(d, sizeof(int32_t) * v.size()));
    CK(cudaMemcpy(*d, v.data(), sizeof(int32_t) * v.size(), cudaMemcpyHostToDevice));
    return 0;
}

int upload_regs(const std::vector<StripReg> &v, StripReg **d)
{
    free_dev(*d);
    *d = nullptr;
    if (v.empty()) return 0;
    CK(cudaMalloc(d, sizeof(StripReg) * v.size()));
    CK(cudaMemcpy(*d, v.data(), sizeof(StripReg) * v.size(), cudaMemcpyHostToDevice));
    return 0;
}

int install_plan(Plan &&pl)
{
    if (!pl.error.empty()) return fail("unsupported mesh: %s", pl.error.c_str());
    if (pl.count != g.nelem) return fail("topology gives %lld local elements but bwk_init was called with %d", (long long)pl.count, g.nelem);
    g.plan = std::move(pl);
    const Plan &P = g.plan;
    const int64_t kq = g.kq;
    if (!g.d_tmp) CK(cudaMalloc(&g.d_tmp, sizeof(double) * g.q_doubles));
    if (!g.d_rsph) CK(cudaMalloc(&g.d_rsph, sizeof(double) * 16 * g.nelem));
    free_dev(g.d_dss);
    CK(cudaMalloc(&g.d_dss, sizeof(ElemDss) * (size_t)g.nelem));
    CK(cudaMemcpy(g.d_dss, P.dss.data(), sizeof(ElemDss) * (size_t)g.nelem, cudaMemcpyHostToDevice));
    if (upload_list(P.boundary_list, &g.d_boundary) || upload_list(P.interior_list, &g.d_interior) || upload_list(P.dependent_list, &g.d_dependent) ||
        upload_list(P.independent_list, &g.d_independent))
        return 1;
    {
        // device strip lists: the full regular strips first (the persistent sweep takes those), the others behind them
        // (face-edge strips with reversed rows, short strips at row ends: sweep2_strip_kernel); both parts keep element order
        const std::vector<int32_t> *S[3] = {&P.strips_all, &P.strips_dependent, &P.strips_independent};
        const std::vector<StripReg> *R[3] = {&P.reg_all, &P.reg_dependent, &P.reg_independent};
        int32_t **ds[3] = {&g.d_strips_all, &g.d_strips_dep, &g.d_strips_indep};
        StripReg **dr[3] = {&g.d_reg_all, &g.d_reg_dep, &g.d_reg_indep};
        static const int max_members = std::getenv("BWK_WSTRIP_MEMBERS") ? std::atoi(std::getenv("BWK_WSTRIP_MEMBERS")) : 8; // tuning knob
        for (int w = 0; w < 3; ++w) {
            // fast strips first, then groups of 4 / 2 / 1 chained fast strips for the warp-strip sweep with 16 / 8 / 4 members per
            // warp (bwk_plan.cpp: strip_is_fast, strips_chain, group_strips)
            const StripGroups G = group_strips(*S[w], *R[w], max_members);
            const std::vector<StripReg> *grp = G.grp;
            g.nfast[w] = G.nfast;
            if (upload_list(G.strips, ds[w]) || upload_regs(G.regs, dr[w])) return 1;
            for (int gi = 0; gi < 3; ++gi) {
                g.ngrp[w][gi] = (int64_t)grp[gi].size() / (WS_NM[gi] / 4);
                if (upload_regs(grp[gi], &g.d_grp[w][gi])) return 1;
            }
        }
    }
    // halo (receive) buffer and send buffer, rows of [KQ][4] doubles
    const int64_t nrecv = P.recv_offset.back();
    free_dev(g.d_halo);
    g.d_halo = nullptr;
    if (nrecv > 0) CK(cudaMalloc(&g.d_halo, sizeof(double) * 4 * kq * nrecv));
    std::vector<int32_t> se, sr;
    g.send_offset.assign((size_t)P.nranks + 1, 0);
    for (int p = 0; p < P.nranks; ++p) {
        for (const RowRef &r : P.send[(size_t)p]) { se.push_back((int32_t)(r.elem - P.first)); sr.push_back(r.row); }
        g.send_offset[(size_t)p + 1] = (int64_t)se.size();
    }
    free_dev(g.d_send);
    g.d_send = nullptr;
    if (!se.empty()) CK(cudaMalloc(&g.d_send, sizeof(double) * 4 * kq * se.size()));
    if (upload_list(se, &g.d_send_elem) || upload_list(sr, &g.d_send_row)) return 1;

    // tensor map of the sweep-1 output; everything else sweep 2 reads goes through plain pointers
    if (make_slab_map(&g.sp.map_tmp, g.d_tmp, g.kq, g.nelem)) return 1;
    {
        const uint64_t dims[3] = {16, (uint64_t)g.kq, (uint64_t)g.nelem};
        const uint64_t str[2] = {128, (uint64_t)g.kq * 128};
        const uint32_t box[3] = {4, TILE_SLABS, 1};
        CUtensorMapL2promotion npromo = CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
        if (const char *e = std::getenv("BWK_NBR_PROMO")) npromo = (CUtensorMapL2promotion)std::atoi(e); // tuning knob: 0 none, 1 64 B, 2 128 B, 3 256 B
        if (make_map(&g.sp.map_rows, g.d_tmp, 3, dims, str, box, CU_TENSOR_MAP_SWIZZLE_32B, npromo)) return 1;
        g.sp.map_halo = g.sp.map_rows;
        // strip ends of sweep 2: the neighbour's slab as {i, j, KQ, elem}; an i-pair (16 bytes, the TMA minimum) of one j
        // holds one value of a W/E edge column or a corner value
        const uint64_t cdims[4] = {4, 4, (uint64_t)g.kq, (uint64_t)g.nelem};
        const uint64_t cstr[3] = {32, 128, (uint64_t)g.kq * 128};
        const uint32_t pbox[4] = {2, 1, TILE_SLABS, 1}, hpbox[3] = {2, TILE_SLABS, 1};
        if (make_map(&g.sp.map_pt, g.d_tmp, 4, cdims, cstr, pbox, CU_TENSOR_MAP_SWIZZLE_NONE, npromo)) return 1;
        if (make_map(&g.sp.map_halo_pt, g.d_tmp, 3, dims, str, hpbox, CU_TENSOR_MAP_SWIZZLE_NONE)) return 1;
        if (nrecv > 0) {
            const uint64_t hdims[3] = {4, (uint64_t)g.kq, (uint64_t)nrecv};
            const uint64_t hstr[2] = {32, (uint64_t)g.kq * 32};
            if (make_map(&g.sp.map_halo, g.d_halo, 3, hdims, hstr, box, CU_TENSOR_MAP_SWIZZLE_32B)) return 1;
            if (make_map(&g.sp.map_halo_pt, g.d_halo, 3, hdims, hstr, hpbox, CU_TENSOR_MAP_SWIZZLE_NONE)) return 1;
        }
    }
    // warp-strip sweep: views of tmp and qtens, and of the received rows
    if (make_field_maps(&g.fm_tmp, g.d_tmp, g.kq, g.nelem) || make_field_maps(&g.fm_q, g.d_q, g.kq, g.nelem)) return 1;
    for (int i = 0; i < 3; ++i) {
        g.map_halo_nm[i] = g.fm_tmp.rows[i]; // (never dereferenced without a halo)
        if (nrecv > 0) {
            const uint64_t hdims[3] = {4, (uint64_t)g.kq, (uint64_t)nrecv}, hstr[2] = {32, (uint64_t)g.kq * 32};
            const uint32_t bh[3] = {4, 32u / (uint32_t)WS_NM[i], (uint32_t)std::min<int64_t>(WS_NM[i], nrecv)};
            if (make_map(&g.map_halo_nm[i], g.d_halo, 3, hdims, hstr, bh, CU_TENSOR_MAP_SWIZZLE_32B)) return 1;
        }
    }
    g.sp.tmp = g.d_tmp;
    g.sp.halo = g.d_halo;
    g.sp.dss = g.d_dss;
    g.sp.rsph = g.d_rsph;
    g.have_topology = true;
    g.have_rsph = false;
    g.rsph_user = false;
    g.e2e_ready = false;
    return 0;
}

// halo exchange of the rows listed in the plan (HOMME bndry_exchangeV): pack on s_comm, then
// grouped ncclSend/ncclRecv with every peer.  Caller orders s_comm after the producers.
int exchange_halo()
{
    const Plan &P = g.plan;
    const int64_t nsend = g.send_offset.empty() ? 0 : g.send_offset.back();
    if (P.nranks == 1) return 0;
    if (!g.comm) return fail("topology spans %d ranks but bwk_comm_init was not called", P.nranks);
    if (nsend > 0) {
        CK(launch_edge_pack(g.src_is_w ? g.d_w : g.d_tmp, g.d_send_elem, g.d_send_row, nsend, g.kq, g.d_send, g.s_comm));
        g.launches += (nsend + 65534) / 65535;
    }
    const size_t row_doubles = (size_t)4 * g.kq;
    NK(g_nccl.GroupStart());
    for (int p = 0; p < P.nranks; ++p) {
        if (p == P.rank) continue;
        const int64_t ns = g.send_offset[(size_t)p + 1] - g.send_offset[(size_t)p];
        const int64_t nr = P.recv_offset[(size_t)p + 1] - P.recv_offset[(size_t)p];
        if (ns > 0) NK(g_nccl.Send(g.d_send + g.send_offset[(size_t)p] * row_doubles, (size_t)ns * row_doubles, ncclDouble, p, g.comm, g.s_comm));
        if (nr > 0) NK(g_nccl.Recv(g.d_halo + P.recv_offset[(size_t)p] * row_doubles, (size_t)nr * row_doubles, ncclDouble, p, g.comm, g.s_comm));
    }
    NK(g_nccl.GroupEnd());
    return 0;
}

// One "produce -> DSS -> consume" pass: produce(first, count, list) writes rspheremp-scaled element-local values to tmp,
// boundary elements first so that their edge rows travel (ncclSend/Recv on s_comm) while the interior is produced; the
// strip kernel then assembles (and, for MODE_SWEEP2, applies the second Laplacian) into `out`.
template <class Produce>
int dss_pass(Produce produce, int consume_mode, const CUtensorMap *out)
{
    const Plan &P = g.plan;
    if (P.nranks == 1) {
        if (produce(0, g.nelem, (const int32_t *)nullptr, g.s_main)) return 1;
        if (g.phase_event >= 0) CK(cudaEventRecord(g.events[g.phase_event], g.s_main));
        return sweep2_list(0, g.s_main, consume_mode, out);
    }
    static const bool trace = std::getenv("BWK_TRACE") != nullptr; // debugging aid: serialised phases with their durations on stderr
    if (trace) {
        auto lap = [&](const char *what, cudaStream_t st, auto &&fn) -> int {
            CK(cudaDeviceSynchronize());
            const auto t0 = std::chrono::steady_clock::now();
            if (fn()) return 1;
            CK(cudaStreamSynchronize(st));
            const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
            std::fprintf(stderr, "[bwk trace rank %d] %-28s %9.1f us\n", P.rank, what, us);
            return 0;
        };
        if (lap("produce(boundary)", g.s_main, [&] { return produce(0, (int)P.boundary_list.size(), (const int32_t *)g.d_boundary, g.s_main); })) return 1;
        if (g.comm && lap("pack + ncclSend/Recv", g.s_comm, [&] { return exchange_halo(); })) return 1;
        if (lap("produce(interior)", g.s_main, [&] { return produce(0, (int)P.interior_list.size(), (const int32_t *)g.d_interior, g.s_main); })) return 1;
        if (lap("consume(independent strips)", g.s_main, [&] { return sweep2_list(2, g.s_main, consume_mode, out); })) return 1;
        if (lap("consume(dependent strips)", g.s_main, [&] { return sweep2_list(1, g.s_main, consume_mode, out); })) return 1;
        std::fprintf(stderr, "[bwk trace rank %d] boundary %zu interior %zu elements; independent %zu dependent %zu strips\n", P.rank,
                     P.boundary_list.size(), P.interior_list.size(), P.strips_independent.size(), P.strips_dependent.size());
        return 0;
    }
    static const bool old_sched = std::getenv("BWK_MULTI_SCHED") && std::atoi(std::getenv("BWK_MULTI_SCHED")) == 0; // A/B knob: round-1 schedule
    if (old_sched) {
        if (produce(0, (int)P.boundary_list.size(), (const int32_t *)g.d_boundary, g.s_main)) return 1;
        CK(cudaEventRecord(g.ev_b, g.s_main));
        CK(cudaStreamWaitEvent(g.s_comm, g.ev_b, 0));
        if (exchange_halo()) return 1;
        CK(cudaEventRecord(g.ev_halo, g.s_comm));
        if (produce(0, (int)P.interior_list.size(), (const int32_t *)g.d_interior, g.s_main)) return 1;
        if (sweep2_list(2, g.s_main, consume_mode, out)) return 1;
        CK(cudaStreamWaitEvent(g.s_main, g.ev_halo, 0));
        if (sweep2_list(1, g.s_main, consume_mode, out)) return 1;
        // the next pass's exchange overwrites the halo buffer: it must not start before these strips have read it
        CK(cudaEventRecord(g.ev_misc, g.s_main));
        CK(cudaStreamWaitEvent(g.s_comm, g.ev_misc, 0));
        return 0;
    }
    // Three streams, no kernel waits for a kernel it does not depend on:
    //   s_comm (highest priority): produce(boundary) -> edge pack -> ncclSend/Recv        -> ev_halo
    //   s_main:                    produce(interior) -> ev_int -> independent strips (after ev_b: they read boundary elements' tmp)
    //   s_dep:                     after ev_int and ev_halo: dependent strips, concurrently with the independent ones
    // The boundary elements' CTAs are dispatched ahead of the interior grid (priority), their tail and the small dependent-strip
    // launch no longer run on a partly empty GPU (round 1: two serialised small launches per step on s_main).
    CK(cudaEventRecord(g.ev_start, g.s_main));
    CK(cudaStreamWaitEvent(g.s_comm, g.ev_start, 0));
    if (produce(0, (int)P.boundary_list.size(), (const int32_t *)g.d_boundary, g.s_comm)) return 1;
    CK(cudaEventRecord(g.ev_b, g.s_comm));
    if (exchange_halo()) return 1;
    CK(cudaEventRecord(g.ev_halo, g.s_comm));
    if (produce(0, (int)P.interior_list.size(), (const int32_t *)g.d_interior, g.s_main)) return 1;
    CK(cudaEventRecord(g.ev_int, g.s_main));
    CK(cudaStreamWaitEvent(g.s_main, g.ev_b, 0));
    if (sweep2_list(2, g.s_main, consume_mode, out)) return 1;
    CK(cudaStreamWaitEvent(g.s_dep, g.ev_int, 0));
    CK(cudaStreamWaitEvent(g.s_dep, g.ev_halo, 0));
    if (sweep2_list(1, g.s_dep, consume_mode, out)) return 1;
    // join; the next pass's exchange overwrites the halo buffer only after ev_start of that pass, i.e. after these strips
    CK(cudaEventRecord(g.ev_misc, g.s_dep));
    CK(cudaStreamWaitEvent(g.s_main, g.ev_misc, 0));
    return 0;
}

int biharmonic_device(int produce_mode = MODE_SWEEP1, const CUtensorMap *out = nullptr, int consume_mode = MODE_SWEEP2)
{
    if (!g.have_metric) return fail("no element metric on the device");
    if (!g.have_topology) return fail("no topology: call bwk_set_topology_* first");
    if (compute_rspheremp()) return 1;
    return dss_pass([&](int first, int count, const int32_t *list, cudaStream_t st) { return sweep(produce_mode, first, count, list, st); }, consume_mode, out);
}

// advance_hypervis_scalar on the device-resident Qdp (d_q); see bwk_b200.h
int hypervis_device(double dt2, double nu_q, int subcycles, int limiter)
{
    if (!g.d_ratio) return fail("no dp / dpdiss on the device: call bwk_upload_hypervis_fields first");
    if (subcycles < 1) return fail("hypervis_subcycle_q must be >= 1");
    if (!sweep1_pre_fits(g.nlev)) return fail("nlev = %d: the per-level factors do not fit in shared memory", g.nlev);
    if (!g.d_w) {
        CK(cudaMalloc(&g.d_w, sizeof(double) * g.q_doubles));
        if (make_slab_map(&g.map_w, g.d_w, g.kq, g.nelem)) return 1;
        const uint64_t dims[3] = {16, (uint64_t)g.kq, (uint64_t)g.nelem}, str[2] = {128, (uint64_t)g.kq * 128};
        const uint32_t box[3] = {4, TILE_SLABS, 1};
        if (make_map(&g.map_w_rows, g.d_w, 3, dims, str, box, CU_TENSOR_MAP_SWIZZLE_32B)) return 1;
        const uint64_t cdims[4] = {4, 4, (uint64_t)g.kq, (uint64_t)g.nelem}, cstr[3] = {32, 128, (uint64_t)g.kq * 128};
        const uint32_t pbox[4] = {2, 1, TILE_SLABS, 1};
        if (make_map(&g.map_w_pt, g.d_w, 4, cdims, cstr, pbox, CU_TENSOR_MAP_SWIZZLE_NONE)) return 1;
        if (make_field_maps(&g.fm_w, g.d_w, g.kq, g.nelem)) return 1;
    }
    g.sp.prescale = g.d_ratio;
    g.sp.nlev = g.nlev;
    g.sp.elem = g.d_elem;
    const double coef = (dt2 / subcycles) * nu_q;
    g.sp.hv_coef = coef;
    g.sp.hv_limiter = limiter;
    static const bool unfused = std::getenv("BWK_HV_UNFUSED") != nullptr; // A/B knob: the 4-launch form with a separate update kernel
    for (int ic = 0; ic < subcycles; ++ic) {
        if (unfused) {
            // Qtens = biharmonic(dpdiss * Qdp / dp) -> work buffer (Qdp itself is still needed)
            if (biharmonic_device(MODE_SWEEP1_PRE, &g.map_w)) return 1;
            // Qdp = DSS(limiter(Qdp * spheremp - dt nu_q Qtens)) * rspheremp
            auto post = [&](int first, int count, const int32_t *list, cudaStream_t st) {
                CK(launch_hypervis_post(g.d_q, g.d_w, g.d_tmp, g.d_elem, g.d_rsph, list, first, count, g.kq, coef, limiter, st));
                g.launches++;
                return 0;
            };
            if (dss_pass(post, MODE_DSS, &g.sp.map_q)) return 1;
            continue;
        }
        // pass 1: sweep 1 on dpdiss*Qdp/dp, then the strip sweep with the update fused in (it TMA-loads the slab of Qdp into
        // the tile's own stage while the second Laplacian runs): work buffer = limiter(Qdp*spheremp - dt nu_q Qtens)*rspheremp
        if (biharmonic_device(MODE_SWEEP1_PRE, &g.map_w, MODE_SWEEP2_POST)) return 1;
        // pass 2: DSS of the work buffer -> Qdp.  Nothing to produce: the halo exchange starts as soon as pass 1 is done and
        // overlaps the independent strips.
        g.src_is_w = true;
        const int rc = dss_pass([](int, int, const int32_t *, cudaStream_t) { return 0; }, MODE_DSS, &g.sp.map_q);
        g.src_is_w = false;
        if (rc) return 1;
    }
    return 0;
}

// biharmonic_wk_dp3d on the device-resident fields (d_q in HOMME's edge-buffer order: T, v, dp3d; the context was made with
// qsize = 4): element-local sweep (-> tmp, rspheremp-scaled, boundary elements first so their rows travel early), DSS strip
// sweep (-> d_q), element-local sweep in place.  See bwk_b200.h
int dp3d_device(int var_coef1, int undamp)
{
    if (g.qsize != 4) return fail("biharmonic_wk_dp3d needs a context made with qsize = 4 (T, v(1), v(2), dp3d per level); this one has qsize = %d", g.qsize);
    if (!g.have_metric) return fail("no element metric on the device");
    if (!g.have_v2c) return fail("no vec_sphere2cart on the device: call bwk_upload_vec_sphere2cart first");
    if (!g.have_topology) return fail("no topology: call bwk_set_topology_* first");
    if (compute_rspheremp()) return 1;
    const double *metric1 = g.d_metric;
    if (!var_coef1) {
        if (!g.d_metric_plain) CK(cudaMalloc(&g.d_metric_plain, sizeof(double) * METRIC_DOUBLES * g.nelem));
        if (!g.have_metric_plain) {
            CK(launch_fold_metric(g.d_elem, g.d_metric_plain, g.nelem, g.s_main, 0));
            g.launches++;
            g.have_metric_plain = true;
        }
        metric1 = g.d_metric_plain;
    }
    auto produce = [&](int first, int count, const int32_t *list, cudaStream_t st) {
        CK(launch_dp3d_laplace(true, g.d_q, g.d_tmp, metric1, g.d_v2c, g.d_elem, g.d_rsph, list, first, count, g.nlev, undamp, st));
        if (count > 0) g.launches++;
        return 0;
    };
    if (dss_pass(produce, MODE_DSS, &g.sp.map_q)) return 1;
    CK(launch_dp3d_laplace(false, g.d_q, g.d_q, g.d_metric, g.d_v2c, g.d_elem, g.d_rsph, nullptr, 0, g.nelem, g.nlev, undamp, g.s_main));
    g.launches++;
    return 0;
}

// chunk boundaries (in elements) for the host<->device pipelines
std::vector<int> make_chunks(int first0, int count)
{
    const double bytes_per_elem = 128.0 * g.kq;
    const double total = bytes_per_elem * count;
    static const double max_chunk_mb = std::getenv("BWK_CHUNK_MB") ? std::atof(std::getenv("BWK_CHUNK_MB")) : 128.0; // tuning knob
    double chunk_bytes = std::min(max_chunk_mb * 1048576.0, std::max(1048576.0, total / 8.0));
    int per = (int)std::max(1.0, std::floor(chunk_bytes / bytes_per_elem));
    std::vector<int> b;
    for (int e = first0; e < first0 + count; e += per) b.push_back(e);
    b.push_back(first0 + count);
    return b;
}

// Schedule of the host-buffer pipeline.  Element chunks are uploaded in ascending order and swept
// (sweep 1) as they arrive.  A strip can run sweep 2 as soon as sweep 1 has produced every element
// it gathers from -- its members and their edge / corner neighbours -- i.e. after upload chunk
// need(strip) = max chunk of those elements; strips that read rows of other ranks wait for the last
// chunk (the halo exchange follows the last sweep 1).  Strips are grouped by need(), each group
// is launched right after the sweep 1 of its chunk and its elements are downloaded behind it, so the
// download of the early elements overlaps the upload of the late ones (PCIe is full duplex).  On the
// cubed sphere only the elements along face edges that touch a later face wait long.
int build_e2e_schedule()
{
    if (g.e2e_ready) return 0;
    const Plan &P = g.plan;
    const std::vector<int32_t> &S = P.strips_all;
    g.e2e_bounds = make_chunks(0, g.nelem);
    const std::vector<int> &b = g.e2e_bounds;
    const int nch = (int)b.size() - 1;
    auto chunk_of = [&](int e) { return (int)(std::upper_bound(b.begin(), b.end(), e) - b.begin()) - 1; };
    std::vector<int> need(S.size());
    for (size_t i = 0; i < S.size(); ++i) {
        const int f = strip_first(S[i]), n = strip_count(S[i]);
        int hi = f + n - 1;
        bool halo = false;
        for (int e = f; e < f + n; ++e) {
            const ElemDss &d = P.dss[(size_t)e];
            for (int k = 0; k < 4; ++k) {
                if (d.edge_src[k] >= 0) { if (d.edge_kind[k] == SRC_HALO) halo = true; else hi = std::max(hi, (int)d.edge_src[k]); }
                if (d.diag_src[k] >= 0) { if (d.diag_kind[k] == SRC_HALO) halo = true; else hi = std::max(hi, (int)d.diag_src[k]); }
            }
        }
        need[i] = halo ? nch - 1 : chunk_of(hi);
    }
    std::vector<int> order(S.size());
    for (size_t i = 0; i < S.size(); ++i) order[i] = (int)i;
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return need[(size_t)x] < need[(size_t)y]; });
    std::vector<int32_t> hs(S.size());
    std::vector<StripReg> hr(S.size());
    g.e2e_group_start.assign((size_t)nch + 1, 0);
    g.e2e_runs.assign((size_t)nch, {});
    for (size_t k = 0; k < order.size(); ++k) {
        const size_t i = (size_t)order[k];
        hs[k] = S[i];
        hr[k] = P.reg_all[i];
        g.e2e_group_start[(size_t)need[i] + 1]++;
        const int f = strip_first(S[i]), n = strip_count(S[i]);
        auto &runs = g.e2e_runs[(size_t)need[i]];
        if (!runs.empty() && runs.back().second == f) runs.back().second = f + n; // strips of a group come in element order
        else runs.push_back({f, f + n});
    }
    for (int c = 0; c < nch; ++c) g.e2e_group_start[(size_t)c + 1] += g.e2e_group_start[(size_t)c];
    free_dev(g.d_e2e_strips);
    free_dev(g.d_e2e_reg);
    g.d_e2e_strips = nullptr;
    g.d_e2e_reg = nullptr;
    if (!S.empty()) {
        CK(cudaMalloc(&g.d_e2e_strips, sizeof(int32_t) * hs.size()));
        CK(cudaMalloc(&g.d_e2e_reg, sizeof(StripReg) * hr.size()));
        CK(cudaMemcpy(g.d_e2e_strips, hs.data(), sizeof(int32_t) * hs.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(g.d_e2e_reg, hr.data(), sizeof(StripReg) * hr.size(), cudaMemcpyHostToDevice));
    }
    g.e2e_ready = true;
    return 0;
}

int ensure_chunk_events(size_t n)
{
    while (g.chunk_up.size() < n) {
        cudaEvent_t a, b;
        CK(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        g.chunk_up.push_back(a);
        g.chunk_done.push_back(b);
    }
    return 0;
}

} // namespace

// =========================================================================================
extern "C" {

const char *bwk_last_error(void) { return g_err.c_str(); }

int bwk_init(int nelemd, int nlev, int qsize)
{
    if (g.inited) return fail("bwk_init called twice without bwk_finalize");
    if (nelemd < 1 || nlev < 1 || qsize < 1) return fail("bad sizes nelemd=%d nlev=%d qsize=%d", nelemd, nlev, qsize);
    if ((int64_t)nlev * qsize > 0x7fffffff / 64) return fail("nlev*qsize too large");
    if (nelemd >= (1 << 28)) return fail("nelemd too large (strip ids are packed into 32 bits)");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev < 1) return fail("no CUDA device available (%s); libbwk_b200 has no CPU fallback", cudaGetErrorString(e));
    cudaDeviceProp prop;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10) return fail("device %d is sm_%d%d; this library is built for sm_100a (B200) only", dev, prop.major, prop.minor);
    g = Ctx();
    if (const char *e = std::getenv("BWK_SWEEP2_TPU")) set_sweep2_tiles_per_unit(std::atoi(e));
    // profiling aid: plan and launch lists of rank r of n on ONE GPU without a communicator ("r,n"; only with BWK_TRACE,
    // halo rows are never filled, results are meaningless)
    if (const char *e = std::getenv("BWK_FAKE_RANKS")) std::sscanf(e, "%d,%d", &g.rank, &g.nranks);
    g.nelem = nelemd;
    g.nlev = nlev;
    g.qsize = qsize;
    g.kq = nlev * qsize;
    g.q_doubles = (size_t)16 * g.kq * nelemd;
    CK(cudaStreamCreateWithFlags(&g.s_main, cudaStreamNonBlocking));
    {
        // the halo pack and the NCCL send/recv kernels must be dispatched as soon as CTA slots free up, not behind the
        // whole interior grid that is already queued on s_main: highest priority for the comm stream
        int prio_lo = 0, prio_hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        CK(cudaStreamCreateWithPriority(&g.s_comm, cudaStreamNonBlocking, prio_hi));
    }
    CK(cudaStreamCreateWithFlags(&g.s_dep, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&g.s_aux, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&g.ev_aux0, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&g.ev_aux1, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&g.s_h2d, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&g.s_d2h, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&g.ev_b, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&g.ev_halo, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&g.ev_misc, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&g.ev_start, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&g.ev_int, cudaEventDisableTiming));
    for (int i = 0; i < NEVENTS; ++i) CK(cudaEventCreate(&g.events[i]));
    CK(cudaMalloc(&g.d_q, sizeof(double) * g.q_doubles));
    CK(cudaMalloc(&g.d_elem, sizeof(double) * ELEM_DOUBLES * nelemd));
    CK(cudaMalloc(&g.d_metric, sizeof(double) * METRIC_DOUBLES * nelemd));
    CK(cudaMalloc(&g.d_partials, sizeof(double) * 2 * NPARTIAL_BLOCKS));
    CK(cudaMalloc(&g.d_work, sizeof(int)));
    CK(cudaMallocHost(&g.h_partials, sizeof(double) * 2 * NPARTIAL_BLOCKS));
    if (make_slab_map(&g.sp.map_q, g.d_q, g.kq, nelemd)) return 1;
    // until a topology is installed map_tmp aliases map_q (never dereferenced by MODE_LAPLACE)
    g.sp.map_tmp = g.sp.map_q;
    g.sp.map_out = g.sp.map_q;
    g.sp.metric = g.d_metric;
    g.sp.kq = g.kq;
    g.inited = true;
    return 0;
}

int bwk_finalize(void)
{
    if (!g.inited) return 0;
    cudaDeviceSynchronize();
    if (g.comm && g_nccl.CommDestroy) g_nccl.CommDestroy(g.comm);
    free_dev(g.d_q); free_dev(g.d_tmp); free_dev(g.d_elem); free_dev(g.d_metric); free_dev(g.d_rsph);
    for (auto &kv : g_tune)
        for (int c = 0; c < 2; ++c)
            for (int e = 0; e < 2; ++e)
                if (kv.second.ev[c][e]) cudaEventDestroy(kv.second.ev[c][e]);
    g_tune.clear();
    free_dev(g.d_work);
    for (auto &row : g.d_grp) for (StripReg *&q : row) { free_dev(q); q = nullptr; }
    free_dev(g.d_w); free_dev(g.d_ratio); free_dev(g.d_v2c); free_dev(g.d_metric_plain);
    free_dev(g.d_save); free_dev(g.d_table); free_dev(g.d_partials); free_dev(g.d_halo); free_dev(g.d_send); free_dev(g.d_dummy);
    free_dev(g.d_dss); free_dev(g.d_boundary); free_dev(g.d_interior); free_dev(g.d_dependent); free_dev(g.d_independent);
    free_dev(g.d_send_elem); free_dev(g.d_send_row); free_dev(g.d_strips_all); free_dev(g.d_strips_dep); free_dev(g.d_strips_indep); free_dev(g.d_reg_all); free_dev(g.d_reg_dep); free_dev(g.d_reg_indep); free_dev(g.d_e2e_strips); free_dev(g.d_e2e_reg);
    if (g.h_partials) cudaFreeHost(g.h_partials);
    for (cudaEvent_t e : g.chunk_up) cudaEventDestroy(e);
    for (cudaEvent_t e : g.chunk_done) cudaEventDestroy(e);
    for (int i = 0; i < NEVENTS; ++i) if (g.events[i]) cudaEventDestroy(g.events[i]);
    if (g.ev_b) cudaEventDestroy(g.ev_b);
    if (g.ev_halo) cudaEventDestroy(g.ev_halo);
    if (g.ev_misc) cudaEventDestroy(g.ev_misc);
    if (g.ev_start) cudaEventDestroy(g.ev_start);
    if (g.ev_int) cudaEventDestroy(g.ev_int);
    if (g.s_dep) cudaStreamDestroy(g.s_dep);
    if (g.s_aux) cudaStreamDestroy(g.s_aux);
    if (g.ev_aux0) cudaEventDestroy(g.ev_aux0);
    if (g.ev_aux1) cudaEventDestroy(g.ev_aux1);
    if (g.s_main) cudaStreamDestroy(g.s_main);
    if (g.s_comm) cudaStreamDestroy(g.s_comm);
    if (g.s_h2d) cudaStreamDestroy(g.s_h2d);
    if (g.s_d2h) cudaStreamDestroy(g.s_d2h);
    g = Ctx();
    return 0;
}

int bwk_upload(const bwk_element_t *elem, const double *qtens, const bwk_derivative_t *deriv)
{
    if (need_init()) return 1;
    if (upload_metric(elem, deriv, g.s_main)) return 1;
    if (qtens) CK(cudaMemcpyAsync(g.d_q, qtens, sizeof(double) * g.q_doubles, cudaMemcpyHostToDevice, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    return 0;
}

int bwk_run(int nets, int nete)
{
    if (need_init() || check_range(nets, nete)) return 1;
    return laplace_range(nets - 1, nete - nets + 1, g.s_main);
}

int bwk_download(double *qtens)
{
    if (need_init()) return 1;
    CK(cudaMemcpyAsync(qtens, g.d_q, sizeof(double) * g.q_doubles, cudaMemcpyDeviceToHost, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    return 0;
}

int bwk_sync(void)
{
    if (need_init()) return 1;
    CK(cudaStreamSynchronize(g.s_main));
    CK(cudaStreamSynchronize(g.s_comm));
    CK(cudaStreamSynchronize(g.s_dep));
    CK(cudaStreamSynchronize(g.s_aux));
    CK(cudaStreamSynchronize(g.s_h2d));
    CK(cudaStreamSynchronize(g.s_d2h));
    return 0;
}

static int e2e_laplace(const bwk_element_t *elem, double *q_host_elem0 /*host address of local element 0*/, const bwk_derivative_t *deriv, int nets,
                       int nete)
{
    if (need_init() || check_range(nets, nete)) return 1;
    if (upload_metric(elem, deriv, g.s_main)) return 1;
    const int first0 = nets - 1, count = nete - nets + 1;
    if (count <= 0) return 0;
    const std::vector<int> b = make_chunks(first0, count);
    const size_t nch = b.size() - 1;
    if (ensure_chunk_events(nch)) return 1;
    const size_t es = (size_t)16 * g.kq; // doubles per element
    for (size_t c = 0; c < nch; ++c) {
        const size_t off = es * (size_t)b[c], n = es * (size_t)(b[c + 1] - b[c]);
        CK(cudaMemcpyAsync(g.d_q + off, q_host_elem0 + off, sizeof(double) * n, cudaMemcpyHostToDevice, g.s_h2d));
        CK(cudaEventRecord(g.chunk_up[c], g.s_h2d));
        CK(cudaStreamWaitEvent(g.s_main, g.chunk_up[c], 0));
        if (laplace_range(b[c], b[c + 1] - b[c], g.s_main)) return 1;
        CK(cudaEventRecord(g.chunk_done[c], g.s_main));
        CK(cudaStreamWaitEvent(g.s_d2h, g.chunk_done[c], 0));
        CK(cudaMemcpyAsync(q_host_elem0 + off, g.d_q + off, sizeof(double) * n, cudaMemcpyDeviceToHost, g.s_d2h));
    }
    CK(cudaStreamSynchronize(g.s_d2h));
    CK(cudaStreamSynchronize(g.s_main));
    return 0;
}

int bwk_biharmonic_wk_scalar(const bwk_element_t *elem, double *qtens, double *grads, const bwk_derivative_t *deriv, const int *nets, const int *nete)
{
    (void)grads; // GPU2's global scratch (F90:529) is not needed by the fused kernel
    if (!elem || !qtens || !deriv || !nets || !nete) return fail("null argument");
    return e2e_laplace(elem, qtens, deriv, *nets, *nete);
}

int bwk_biharmonic_wk_scalar_nets(const bwk_element_t *elem, double *qtens_nets, const bwk_derivative_t *deriv, const int *nets, const int *nete)
{
    if (!elem || !qtens_nets || !deriv || !nets || !nete) return fail("null argument");
    if (need_init()) return 1;
    // qtens(np,np,nlev,qsize,nets:nete): the dummy's first element is element nets (F90:188)
    return e2e_laplace(elem, qtens_nets - (size_t)16 * g.kq * (size_t)(*nets - 1), deriv, *nets, *nete);
}

int bwk_biharmonic_wk_scalar_full(const bwk_element_t *elem, double *qtens, const bwk_derivative_t *deriv)
{
    if (need_init()) return 1;
    if (!elem || !qtens || !deriv) return fail("null argument");
    if (!g.have_topology) return fail("no topology: call bwk_set_topology_* first");
    if (upload_metric(elem, deriv, g.s_main)) return 1;
    if (compute_rspheremp()) return 1;
    if (build_e2e_schedule()) return 1;
    const std::vector<int> &b = g.e2e_bounds;
    const size_t nch = b.size() - 1;
    if (ensure_chunk_events(nch)) return 1;
    const size_t es = (size_t)16 * g.kq;
    for (size_t c = 0; c < nch; ++c) {
        // upload chunk c, sweep 1 on it as soon as it has landed
        const size_t off = es * (size_t)b[c], n = es * (size_t)(b[c + 1] - b[c]);
        CK(cudaMemcpyAsync(g.d_q + off, qtens + off, sizeof(double) * n, cudaMemcpyHostToDevice, g.s_h2d));
        CK(cudaEventRecord(g.chunk_up[c], g.s_h2d));
        CK(cudaStreamWaitEvent(g.s_main, g.chunk_up[c], 0));
        if (sweep(MODE_SWEEP1, b[c], b[c + 1] - b[c], nullptr, g.s_main)) return 1;
        if (c + 1 == nch && g.plan.nranks > 1) { // every edge row exists now: exchange the halo (bndry_exchangeV)
            CK(cudaEventRecord(g.ev_b, g.s_main));
            CK(cudaStreamWaitEvent(g.s_comm, g.ev_b, 0));
            if (exchange_halo()) return 1;
            CK(cudaEventRecord(g.ev_halo, g.s_comm));
            CK(cudaStreamWaitEvent(g.s_main, g.ev_halo, 0));
        }
        // sweep 2 of the strips whose inputs are complete with this chunk, download behind it
        const int s0 = g.e2e_group_start[c], s1 = g.e2e_group_start[c + 1];
        if (s1 > s0) {
            if (sweep2(g.d_e2e_strips + s0, g.d_e2e_reg + s0, s1 - s0, g.s_main)) return 1;
            CK(cudaEventRecord(g.chunk_done[c], g.s_main));
            CK(cudaStreamWaitEvent(g.s_d2h, g.chunk_done[c], 0));
            for (const std::pair<int, int> &r : g.e2e_runs[c]) {
                const size_t roff = es * (size_t)r.first, rn = es * (size_t)(r.second - r.first);
                CK(cudaMemcpyAsync(qtens + roff, g.d_q + roff, sizeof(double) * rn, cudaMemcpyDeviceToHost, g.s_d2h));
            }
        }
    }
    CK(cudaStreamSynchronize(g.s_d2h));
    CK(cudaStreamSynchronize(g.s_main));
    return 0;
}

int bwk_host_alloc(void **ptr, uint64_t bytes)
{
    CK(cudaMallocHost(ptr, bytes));
    return 0;
}
int bwk_host_free(void *ptr)
{
    CK(cudaFreeHost(ptr));
    return 0;
}
int bwk_host_register(void *ptr, uint64_t bytes)
{
    CK(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
    return 0;
}
int bwk_host_unregister(void *ptr)
{
    CK(cudaHostUnregister(ptr));
    return 0;
}

int bwk_set_variant(int variant)
{
    if (variant != 0 && variant != 1) return fail("unknown variant %d", variant);
    g.variant = variant;
    return 0;
}
int64_t bwk_launch_count(void) { return g.launches; }
int bwk_get_stream(uint64_t *stream)
{
    if (need_init()) return 1;
    *stream = (uint64_t)(uintptr_t)g.s_main;
    return 0;
}
int bwk_get_device_qtens(uint64_t *ptr)
{
    if (need_init()) return 1;
    *ptr = (uint64_t)(uintptr_t)g.d_q;
    return 0;
}

// ------------------------------------------------------------------------------ harness
int bwk_initialize_data_device(int64_t nelemd_global, int64_t elem_first, int seed_mode)
{
    if (need_init()) return 1;
    if (elem_first < 0 || elem_first + g.nelem > nelemd_global) return fail("elements [%lld,%lld) outside global 0..%lld", (long long)elem_first,
                                                                            (long long)(elem_first + g.nelem), (long long)nelemd_global);
    static const std::vector<double> table = lcg_table();
    if (!g.d_table) { // once per context: a 1 MB copy from pageable memory would put a host-side wait into every call
        CK(cudaMalloc(&g.d_table, sizeof(double) * 131072));
        CK(cudaMemcpyAsync(g.d_table, table.data(), sizeof(double) * 131072, cudaMemcpyHostToDevice, g.s_main));
        CK(cudaStreamSynchronize(g.s_main));
    }
    double dvv[16];
    for (int t = 0; t < 16; ++t) {
        if (seed_mode == 0) dvv[t] = table[(size_t)t];
        else {
            uint64_t x = (uint64_t)t + 0x9E3779B97F4A7C15ull;
            x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
            x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
            x ^= x >> 31;
            dvv[t] = (double)(x >> 11) * 0x1.0p-53;
        }
    }
    CK(set_dvv(dvv, g.s_main));
    CK(launch_init_data(g.d_table, nelemd_global, elem_first, g.nelem, g.kq, seed_mode, g.d_elem, g.d_q, g.s_main));
    CK(launch_fold_metric(g.d_elem, g.d_metric, g.nelem, g.s_main));
    g.launches += 2;
    g.have_metric = true;
    const bool same = g.gen_nelem_global == nelemd_global && g.gen_elem_first == elem_first && g.gen_seed_mode == seed_mode &&
                      (int64_t)g.h_spheremp.size() == (int64_t)16 * g.nelem;
    if (!same) {
        // keep the local spheremp on the host for rspheremp
        std::vector<double> he((size_t)ELEM_DOUBLES * g.nelem);
        CK(cudaMemcpyAsync(he.data(), g.d_elem, sizeof(double) * he.size(), cudaMemcpyDeviceToHost, g.s_main));
        CK(cudaStreamSynchronize(g.s_main));
        g.h_spheremp.resize((size_t)16 * g.nelem);
        for (int64_t ie = 0; ie < g.nelem; ++ie) std::memcpy(&g.h_spheremp[(size_t)16 * ie], &he[(size_t)(ie * ELEM_DOUBLES + 64)], 16 * sizeof(double));
        if (!g.rsph_user) g.have_rsph = false; // a caller-supplied rspheremp stays in force until the topology changes
        g.gen_nelem_global = nelemd_global;
        g.gen_elem_first = elem_first;
        g.gen_seed_mode = seed_mode;
    }
    return 0;
}

int bwk_save_qtens(void)
{
    if (need_init()) return 1;
    if (!g.d_save) CK(cudaMalloc(&g.d_save, sizeof(double) * g.q_doubles));
    CK(cudaMemcpyAsync(g.d_save, g.d_q, sizeof(double) * g.q_doubles, cudaMemcpyDeviceToDevice, g.s_main));
    return 0;
}

int bwk_compute_l2norm(double *l2)
{
    if (need_init()) return 1;
    if (!g.d_save) return fail("bwk_save_qtens has not been called");
    CK(launch_l2_partials(g.d_q, g.d_save, (int64_t)g.q_doubles, g.d_partials, NPARTIAL_BLOCKS, g.s_main));
    g.launches++;
    CK(cudaMemcpyAsync(g.h_partials, g.d_partials, sizeof(double) * 2 * NPARTIAL_BLOCKS, cudaMemcpyDeviceToHost, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    double num = 0, den = 0;
    for (int i = 0; i < NPARTIAL_BLOCKS; ++i) { num += g.h_partials[i]; den += g.h_partials[NPARTIAL_BLOCKS + i]; }
    *l2 = std::sqrt(num / den);
    return 0;
}

int bwk_checksum(double *sum, double *sumsq)
{
    if (need_init()) return 1;
    CK(launch_sum_partials(g.d_q, (int64_t)g.q_doubles, g.d_partials, NPARTIAL_BLOCKS, g.s_main));
    g.launches++;
    CK(cudaMemcpyAsync(g.h_partials, g.d_partials, sizeof(double) * 2 * NPARTIAL_BLOCKS, cudaMemcpyDeviceToHost, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    double a = 0, b = 0;
    for (int i = 0; i < NPARTIAL_BLOCKS; ++i) { a += g.h_partials[i]; b += g.h_partials[NPARTIAL_BLOCKS + i]; }
    *sum = a;
    *sumsq = b;
    return 0;
}

int bwk_checksum_bits(int64_t elem_first_global, uint64_t *hash)
{
    if (need_init()) return 1;
    if (elem_first_global < 0 || !hash) return fail("bad arguments");
    static_assert(sizeof(unsigned long long) == sizeof(double), "partials buffer is reused");
    unsigned long long *dp = reinterpret_cast<unsigned long long *>(g.d_partials), *hp = reinterpret_cast<unsigned long long *>(g.h_partials);
    CK(launch_bits_checksum(g.d_q, (int64_t)g.q_doubles, (uint64_t)elem_first_global * 16ull * (uint64_t)g.kq, dp, NPARTIAL_BLOCKS, g.s_main));
    g.launches++;
    CK(cudaMemcpyAsync(hp, dp, sizeof(unsigned long long) * NPARTIAL_BLOCKS, cudaMemcpyDeviceToHost, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    unsigned long long a = 0;
    for (int i = 0; i < NPARTIAL_BLOCKS; ++i) a += hp[i];
    *hash = a;
    return 0;
}

int bwk_periodicity_defect(int64_t period_elems, double *max_abs_diff)
{
    if (need_init()) return 1;
    if (period_elems < 1) return fail("bad period");
    CK(launch_period_defect(g.d_q, (int64_t)g.q_doubles, period_elems * 16 * g.kq, g.d_partials, NPARTIAL_BLOCKS, g.s_main));
    g.launches++;
    CK(cudaMemcpyAsync(g.h_partials, g.d_partials, sizeof(double) * NPARTIAL_BLOCKS, cudaMemcpyDeviceToHost, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    double m = 0;
    for (int i = 0; i < NPARTIAL_BLOCKS; ++i) m = std::max(m, g.h_partials[i]);
    *max_abs_diff = m;
    return 0;
}

int bwk_download_elements(int64_t first, int64_t count, double *dst)
{
    if (need_init()) return 1;
    if (first < 0 || count < 0 || first + count > g.nelem) return fail("bad element range");
    const size_t es = (size_t)16 * g.kq;
    CK(cudaMemcpyAsync(dst, g.d_q + es * (size_t)first, sizeof(double) * es * (size_t)count, cudaMemcpyDeviceToHost, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    return 0;
}

int bwk_upload_elements(int64_t first, int64_t count, const double *src)
{
    if (need_init()) return 1;
    if (first < 0 || count < 0 || first + count > g.nelem) return fail("bad element range");
    const size_t es = (size_t)16 * g.kq;
    CK(cudaMemcpyAsync(g.d_q + es * (size_t)first, src, sizeof(double) * es * (size_t)count, cudaMemcpyHostToDevice, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    return 0;
}

// ------------------------------------------------------------------------------ full biharmonic
int bwk_cubed_sphere_gdof(int ne, int64_t *gdof)
{
    if (ne < 1 || !gdof) return fail("bad arguments");
    cubed_sphere_gdof(ne, gdof);
    return 0;
}

int bwk_set_topology_gdof(int64_t nelem_global, const int64_t *gdof, int64_t elem_first)
{
    if (need_init()) return 1;
    int64_t f, c;
    partition(nelem_global, g.rank, g.nranks, &f, &c);
    if (f != elem_first || c != g.nelem)
        return fail("rank %d of %d owns elements [%lld,%lld) of %lld, but the context holds %d elements from %lld", g.rank, g.nranks, (long long)f,
                    (long long)(f + c), (long long)nelem_global, g.nelem, (long long)elem_first);
    return install_plan(build_plan(nelem_global, gdof, g.rank, g.nranks));
}

int bwk_set_topology_cubed_sphere(int ne, int64_t elem_first)
{
    if (ne < 1) return fail("bad ne");
    const int64_t n = (int64_t)6 * ne * ne;
    std::vector<int64_t> gdof((size_t)16 * n);
    cubed_sphere_gdof(ne, gdof.data());
    return bwk_set_topology_gdof(n, gdof.data(), elem_first);
}

int bwk_set_rspheremp(const double *rs)
{
    if (need_init()) return 1;
    if (!g.d_rsph) CK(cudaMalloc(&g.d_rsph, sizeof(double) * 16 * g.nelem));
    CK(cudaMemcpy(g.d_rsph, rs, sizeof(double) * 16 * g.nelem, cudaMemcpyHostToDevice));
    g.sp.rsph = g.d_rsph;
    g.have_rsph = true;
    g.rsph_user = true;
    return 0;
}

int bwk_get_rspheremp(double *rs)
{
    if (need_init()) return 1;
    if (compute_rspheremp()) return 1;
    CK(cudaMemcpy(rs, g.d_rsph, sizeof(double) * 16 * g.nelem, cudaMemcpyDeviceToHost));
    return 0;
}

int bwk_run_biharmonic(void)
{
    if (need_init()) return 1;
    return biharmonic_device();
}

int bwk_measure_fp64_peak(double *tflops)
{
    if (need_init()) return 1;
    if (!g.d_dummy) CK(cudaMalloc(&g.d_dummy, sizeof(double) * 2 * 1024));
    const int blocks = 148 * 32;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(launch_dfma_peak(g.d_dummy, blocks, g.s_main)); // warm-up
    CK(cudaEventRecord(e0, g.s_main));
    for (int r = 0; r < 5; ++r) CK(launch_dfma_peak(g.d_dummy, blocks, g.s_main));
    CK(cudaEventRecord(e1, g.s_main));
    CK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    g.launches += 6;
    *tflops = 5.0 * blocks * 256.0 * 8.0 * 4096.0 * 2.0 / (ms * 1e-3) / 1e12;
    return 0;
}

int bwk_upload_hypervis_fields(const double *dp, const double *dpdiss)
{
    if (need_init()) return 1;
    if (!dp || !dpdiss) return fail("bwk_upload_hypervis_fields: dp and dpdiss are required");
    const size_t n = (size_t)16 * g.nlev * g.nelem;
    if (!g.d_ratio) CK(cudaMalloc(&g.d_ratio, sizeof(double) * n));
    double *d_den = nullptr;
    CK(cudaMalloc(&d_den, sizeof(double) * n));
    cudaError_t e = cudaMemcpyAsync(g.d_ratio, dpdiss, sizeof(double) * n, cudaMemcpyHostToDevice, g.s_main);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_den, dp, sizeof(double) * n, cudaMemcpyHostToDevice, g.s_main);
    if (e == cudaSuccess) e = launch_ratio(g.d_ratio, d_den, g.d_ratio, (int64_t)n, g.s_main);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g.s_main);
    cudaFree(d_den);
    CK(e);
    g.launches++;
    return 0;
}

int bwk_run_advance_hypervis(double dt2, double nu_q, int hypervis_subcycle_q, int limiter)
{
    if (need_init()) return 1;
    if (!g.have_metric) return fail("no element metric on the device");
    if (!g.have_topology) return fail("no topology: call bwk_set_topology_* first");
    if (compute_rspheremp()) return 1;
    return hypervis_device(dt2, nu_q, hypervis_subcycle_q, limiter);
}

int bwk_advance_hypervis_scalar(const bwk_element_t *elem, double *Qdp, const double *dp, const double *dpdiss,
                                const bwk_derivative_t *deriv, double dt2, double nu_q, int hypervis_subcycle_q, int limiter)
{
    if (need_init()) return 1;
    if (!Qdp) return fail("bwk_advance_hypervis_scalar: Qdp is required");
    if (!g.have_topology) return fail("no topology: call bwk_set_topology_* first");
    if (bwk_upload(elem, Qdp, deriv)) return 1;
    if (dp || dpdiss) { if (bwk_upload_hypervis_fields(dp, dpdiss)) return 1; }
    if (bwk_run_advance_hypervis(dt2, nu_q, hypervis_subcycle_q, limiter)) return 1;
    return bwk_download(Qdp);
}

int bwk_upload_vec_sphere2cart(const double *vec_sphere2cart)
{
    if (need_init()) return 1;
    if (!vec_sphere2cart) return fail("bwk_upload_vec_sphere2cart: the array is required");
    if (!g.d_v2c) CK(cudaMalloc(&g.d_v2c, sizeof(double) * 96 * g.nelem));
    CK(cudaMemcpyAsync(g.d_v2c, vec_sphere2cart, sizeof(double) * 96 * g.nelem, cudaMemcpyHostToDevice, g.s_main));
    CK(cudaStreamSynchronize(g.s_main));
    g.have_v2c = true;
    return 0;
}

int bwk_run_biharmonic_dp3d(int var_coef1, int undamprrcart)
{
    if (need_init()) return 1;
    return dp3d_device(var_coef1, undamprrcart);
}

// host arrays in HOMME's layouts <-> the fused device field (per element: T | v | dp3d, 4*nlev slabs)
static int dp3d_copy(double *dptens, double *ttens, double *vtens, bool to_device)
{
    const size_t lev = (size_t)g.nlev * 128, pitch = 4 * lev;
    char *base = reinterpret_cast<char *>(g.d_q);
    struct Part { double *host; size_t off, width; } parts[3] = {{ttens, 0, lev}, {vtens, lev, 2 * lev}, {dptens, 3 * lev, lev}};
    for (const Part &p : parts) {
        if (to_device) CK(cudaMemcpy2DAsync(base + p.off, pitch, p.host, p.width, p.width, (size_t)g.nelem, cudaMemcpyHostToDevice, g.s_main));
        else CK(cudaMemcpy2DAsync(p.host, p.width, base + p.off, pitch, p.width, (size_t)g.nelem, cudaMemcpyDeviceToHost, g.s_main));
    }
    return 0;
}

int bwk_biharmonic_wk_dp3d(const bwk_element_t *elem, const double *vec_sphere2cart, double *dptens, double *ttens, double *vtens,
                           const bwk_derivative_t *deriv, int var_coef1, int undamprrcart)
{
    if (need_init()) return 1;
    if (!dptens || !ttens || !vtens) return fail("bwk_biharmonic_wk_dp3d: dptens, ttens and vtens are required");
    if (g.qsize != 4) return fail("biharmonic_wk_dp3d needs a context made with qsize = 4; this one has qsize = %d", g.qsize);
    if (!g.have_topology) return fail("no topology: call bwk_set_topology_* first");
    if (upload_metric(elem, deriv, g.s_main)) return 1;
    if (vec_sphere2cart && bwk_upload_vec_sphere2cart(vec_sphere2cart)) return 1;
    if (dp3d_copy(dptens, ttens, vtens, true)) return 1;
    if (dp3d_device(var_coef1, undamprrcart)) return 1;
    if (dp3d_copy(dptens, ttens, vtens, false)) return 1;
    CK(cudaStreamSynchronize(g.s_main));
    return 0;
}

int bwk_run_biharmonic_phase(int phase)
{
    if (need_init()) return 1;
    if (!g.have_metric) return fail("no element metric on the device");
    if (!g.have_topology) return fail("no topology: call bwk_set_topology_* first");
    if (g.plan.nranks != 1) return fail("bwk_run_biharmonic_phase is a single-rank profiling hook");
    if (compute_rspheremp()) return 1;
    if (phase == 1) return sweep(MODE_SWEEP1, 0, g.nelem, nullptr, g.s_main);
    if (phase == 2) return sweep2_list(0, g.s_main);
    return fail("phase must be 1 or 2");
}

// ------------------------------------------------------------------------------ several GPUs
int bwk_comm_unique_id(void *id_bytes)
{
    if (!g_nccl.load()) return fail("cannot load libnccl.so.2: %s", dlerror());
    static_assert(sizeof(ncclUniqueId) == BWK_UNIQUE_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    NK(g_nccl.GetUniqueId(&id));
    std::memcpy(id_bytes, &id, sizeof id);
    return 0;
}

int bwk_comm_init(int rank, int nranks, const void *id_bytes)
{
    if (need_init()) return 1;
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail("bad rank %d of %d", rank, nranks);
    if (g.have_topology) return fail("bwk_comm_init must precede bwk_set_topology_*");
    g.rank = rank;
    g.nranks = nranks;
    if (nranks == 1) return 0;
    if (!g_nccl.load()) return fail("cannot load libnccl.so.2: %s", dlerror());
    ncclUniqueId id;
    std::memcpy(&id, id_bytes, sizeof id);
    NK(g_nccl.CommInitRank(&g.comm, nranks, id, rank));
    return 0;
}

// device-side barrier on the library's stream: every rank exchanges 8 bytes with every other rank, so work enqueued
// after it starts within one NCCL latency on all ranks (a host barrier leaves hundreds of microseconds of skew)
int bwk_comm_barrier(void)
{
    if (need_init()) return 1;
    if (g.nranks == 1 || !g.comm) return 0;
    if (!g.d_dummy) CK(cudaMalloc(&g.d_dummy, sizeof(double) * 2 * 1024));
    NK(g_nccl.GroupStart());
    for (int p = 0; p < g.nranks; ++p) {
        if (p == g.rank) continue;
        NK(g_nccl.Send(g.d_dummy + g.rank, 1, ncclDouble, p, g.comm, g.s_main));
        NK(g_nccl.Recv(g.d_dummy + 1024 + p, 1, ncclDouble, p, g.comm, g.s_main));
    }
    NK(g_nccl.GroupEnd());
    return 0;
}

int bwk_comm_finalize(void)
{
    if (g.comm) {
        cudaDeviceSynchronize();
        NK(g_nccl.CommDestroy(g.comm));
        g.comm = nullptr;
    }
    g.rank = 0;
    g.nranks = 1;
    return 0;
}

int bwk_partition(int64_t nelem_global, int rank, int nranks, int64_t *first, int64_t *count)
{
    if (nranks < 1 || rank < 0 || rank >= nranks || nelem_global < 0) return fail("bad arguments");
    partition(nelem_global, rank, nranks, first, count);
    return 0;
}

// ------------------------------------------------------------------------------ events (timing on the library stream)
int bwk_event_record(int slot)
{
    if (need_init()) return 1;
    if (slot < 0 || slot >= NEVENTS) return fail("event slot out of range");
    CK(cudaEventRecord(g.events[slot], g.s_main));
    return 0;
}
int bwk_set_phase_event(int slot)
{
    if (need_init()) return 1;
    if (slot >= NEVENTS) return fail("event slot out of range");
    g.phase_event = slot < 0 ? -1 : slot;
    return 0;
}
int bwk_event_elapsed_ms(int a, int b, float *ms)
{
    if (need_init()) return 1;
    if (a < 0 || a >= NEVENTS || b < 0 || b >= NEVENTS) return fail("event slot out of range");
    CK(cudaEventSynchronize(g.events[b]));
    CK(cudaEventElapsedTime(ms, g.events[a], g.events[b]));
    return 0;
}

// ------------------------------------------------------------------------------ plan introspection (host only)
struct bwk_plan {
    Plan p;
};

bwk_plan_t *bwk_plan_create(int64_t nelem_global, const int64_t *gdof, int rank, int nranks)
{
    if (nelem_global < 1 || !gdof || nranks < 1 || rank < 0 || rank >= nranks) { fail("bad arguments"); return nullptr; }
    bwk_plan *h = new bwk_plan;
    h->p = build_plan(nelem_global, gdof, rank, nranks);
    if (!h->p.error.empty()) {
        fail("unsupported mesh: %s", h->p.error.c_str());
        delete h;
        return nullptr;
    }
    return h;
}
void bwk_plan_destroy(bwk_plan_t *h) { delete h; }
int64_t bwk_plan_unique_nodes(const bwk_plan_t *h) { return h->p.unique_nodes; }
int64_t bwk_plan_local_first(const bwk_plan_t *h) { return h->p.first; }
int64_t bwk_plan_local_count(const bwk_plan_t *h) { return h->p.count; }
int64_t bwk_plan_recv_rows(const bwk_plan_t *h, int peer) { return (int64_t)h->p.recv[(size_t)peer].size(); }
int64_t bwk_plan_send_rows(const bwk_plan_t *h, int peer) { return (int64_t)h->p.send[(size_t)peer].size(); }
static int copy_list(const std::vector<RowRef> &v, int64_t *elem, int32_t *row)
{
    for (size_t i = 0; i < v.size(); ++i) { elem[i] = v[i].elem; row[i] = v[i].row; }
    return 0;
}
int bwk_plan_send_list(const bwk_plan_t *h, int peer, int64_t *elem, int32_t *row) { return copy_list(h->p.send[(size_t)peer], elem, row); }
int bwk_plan_recv_list(const bwk_plan_t *h, int peer, int64_t *elem, int32_t *row) { return copy_list(h->p.recv[(size_t)peer], elem, row); }
int64_t bwk_plan_boundary_elems(const bwk_plan_t *h) { return (int64_t)h->p.boundary_list.size(); }
int64_t bwk_plan_dependent_elems(const bwk_plan_t *h) { return (int64_t)h->p.dependent_list.size(); }
static const std::vector<int32_t> *strip_vec(const bwk_plan_t *h, int which)
{
    return which == 0 ? &h->p.strips_all : which == 1 ? &h->p.strips_dependent : which == 2 ? &h->p.strips_independent : nullptr;
}
int64_t bwk_plan_strip_count(const bwk_plan_t *h, int which)
{
    const std::vector<int32_t> *v = strip_vec(h, which);
    return v ? (int64_t)v->size() : -1;
}
int bwk_plan_strip_list(const bwk_plan_t *h, int which, int32_t *first, int32_t *count, int32_t *regular)
{
    const std::vector<int32_t> *v = strip_vec(h, which);
    if (!v) return fail("which must be 0, 1 or 2");
    const std::vector<StripReg> &r = which == 0 ? h->p.reg_all : which == 1 ? h->p.reg_dependent : h->p.reg_independent;
    for (size_t i = 0; i < v->size(); ++i) {
        first[i] = strip_first((*v)[i]);
        count[i] = strip_count((*v)[i]);
        if (regular) regular[i] = r[i].regular;
    }
    return 0;
}
int64_t bwk_plan_strip_groups(const bwk_plan_t *h, int which, int max_members, int32_t *first, int32_t *nstrips)
{
    const std::vector<int32_t> *v = strip_vec(h, which);
    if (!v) return -1;
    const std::vector<StripReg> &r = which == 0 ? h->p.reg_all : which == 1 ? h->p.reg_dependent : h->p.reg_independent;
    const StripGroups G = group_strips(*v, r, max_members);
    int64_t n = 0;
    for (int gi = 2; gi >= 0; --gi) { // the order the sweep launches them in does not matter; report the widest groups first
        const int per = 1 << gi;
        for (size_t i = 0; i + per <= G.grp[gi].size(); i += (size_t)per, ++n)
            if (first) { first[n] = G.grp[gi][i].first; nstrips[n] = per; }
    }
    for (size_t i = (size_t)G.nfast; i < G.strips.size(); ++i, ++n)
        if (first) { first[n] = strip_first(G.strips[i]); nstrips[n] = 0; }
    return n;
}
int bwk_plan_apply_host(const bwk_plan_t *h, const double *field, double *out)
{
    for (int64_t le = 0; le < h->p.count; ++le) gather_elem_host(h->p, le, field, out + 16 * le);
    return 0;
}

} // extern "C"

// ------------------------------------------------------------------------------ allocation-free single Laplacian
// The sweep reads the folded metric ([nelemd][64] doubles, written by fold_metric_kernel from elem) as scratch.  Here that
// scratch is a file-scope device array instead of a cudaMalloc'd buffer: every GPU that loads the library holds it until exit.
// Its name below is only the host-side stand-in of the array; the address a kernel can use comes from cudaGetSymbolAddress,
// asked once per device with that device current.
static __device__ __align__(256) double g_noalloc_metric[(size_t)BWK_NOALLOC_MAX_NELEMD * METRIC_DOUBLES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
double *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

int noalloc_scratch(double **ptr)
{
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail("device ordinal %d outside 0..%d", dev, NOALLOC_MAX_DEVICES - 1);
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_metric));
        if (!p) return fail("cudaGetSymbolAddress returned a null address for the scratch array");
        g_noalloc_ptr[dev] = static_cast<double *>(p);
    }
    *ptr = g_noalloc_ptr[dev];
    return 0;
}
} // namespace

extern "C" int bwk_biharmonic_wk_scalar_noalloc(const bwk_element_t *d_elem, double *d_qtens, const bwk_derivative_t *deriv, int nelemd, int nlev,
                                                int qsize, uint64_t stream)
{
    if (nelemd < 1 || nlev < 1 || qsize < 1) return fail("bad sizes nelemd=%d nlev=%d qsize=%d", nelemd, nlev, qsize);
    if (nelemd > BWK_NOALLOC_MAX_NELEMD)
        return fail("bwk_biharmonic_wk_scalar_noalloc: nelemd=%d exceeds BWK_NOALLOC_MAX_NELEMD=%d (the static scratch array holds the folded "
                    "metric of that many elements); use bwk_init/bwk_upload/bwk_run for larger problems",
                    nelemd, BWK_NOALLOC_MAX_NELEMD);
    if ((int64_t)nlev * qsize > 0x7fffffff / 64) return fail("nlev*qsize too large");
    if (!d_elem || !d_qtens || !deriv) return fail("bwk_biharmonic_wk_scalar_noalloc: null pointer argument");
    if ((reinterpret_cast<uintptr_t>(d_qtens) & 127) != 0) return fail("bwk_biharmonic_wk_scalar_noalloc: d_qtens must be 128-byte aligned");
    double *scratch = nullptr;
    if (noalloc_scratch(&scratch)) return 1;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(static_cast<uintptr_t>(stream));
    const int kq = nlev * qsize;
    SweepParams p{};
    if (make_slab_map(&p.map_q, d_qtens, kq, nelemd)) return 1;
    p.map_tmp = p.map_q;
    p.map_out = p.map_q;
    p.metric = scratch;
    p.kq = kq;
    const SweepGeometry geo = choose_geometry(kq, nelemd, MODE_LAPLACE);
    p.tiles_per_unit = geo.tiles_per_unit;
    p.units = geo.units;
    p.elem_first = 0;
    p.elem_count = nelemd;
    CK(set_dvv(&deriv->Dvv[0][0], st));
    // every call rewrites the nelemd*64 doubles it reads: nothing in the array has to start at zero, and what an earlier
    // call left beyond nelemd*64 is never read
    CK(launch_fold_metric(reinterpret_cast<const double *>(d_elem), scratch, nelemd, st));
    CK(launch_sweep(MODE_LAPLACE, p, st));
    return 0;
}
