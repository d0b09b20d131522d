// gsb200_api.cu -- the C ABI of include/gsb200.h: handle, streams, ping-pong buffers, step
// orchestration (interior / shell overlap with the halo exchange), NCCL and in-process
// transports.  Kernels live in gs_stencil.cuh / gs_kernels.cu.
#include <dlfcn.h>
#include <nccl.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/gsb200.h"
#include "gs_kernels.h"

using namespace gsb;

// ------------------------------------------------------------------ errors
static thread_local std::string g_err;

// gs_noalloc.cu reports through the same thread-local message
namespace gsb { void set_last_error(const char* msg) { g_err = msg ? msg : ""; } }

static int fail(const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

#define CK(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess)                                                                        \
            return fail("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));        \
    } while (0)

// Makes `dev` current for the duration of a call and restores the caller's device afterwards (the host
// process -- Julia with CUDA.jl, Python with torch -- has its own notion of the current device).
struct DeviceGuard {
    int prev = -1;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int dev) {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};

// ------------------------------------------------------------------ NCCL, loaded lazily
// dlopen instead of a link-time dependency: a process that already holds an NCCL (PyTorch's
// bundled one) must keep exactly that copy.
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;

static std::mutex g_nccl_mu;

static int nccl_load() {
    std::lock_guard<std::mutex> lock(g_nccl_mu);
    if (g_nccl.lib) return 0;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return fail("cannot load libnccl.so.2: %s", dlerror());
#define SYM(field, name)                                                       \
    *(void**)(&g_nccl.field) = dlsym(lib, name);                               \
    if (!g_nccl.field) return fail("libnccl: missing symbol %s", name);
    SYM(GetUniqueId, "ncclGetUniqueId")
    SYM(CommInitRank, "ncclCommInitRank")
    SYM(CommDestroy, "ncclCommDestroy")
    SYM(Send, "ncclSend")
    SYM(Recv, "ncclRecv")
    SYM(GroupStart, "ncclGroupStart")
    SYM(GroupEnd, "ncclGroupEnd")
    SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    g_nccl.lib = lib;
    return 0;
}

#define NK(call)                                                                                      \
    do {                                                                                              \
        ncclResult_t r_ = (call);                                                                     \
        if (r_ != ncclSuccess)                                                                        \
            return fail("%s:%d: %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_));     \
    } while (0)

// ------------------------------------------------------------------ handle
enum : int { TR_NONE = 0, TR_LOCAL = 1, TR_NCCL = 2 };
enum : int { HALO_PIPELINE = 0, HALO_SPLIT = 1 };
#define GSB_MAX_CHUNKS 16

struct gsb_handle {
    gsb_config cfg;
    Layout L;
    size_t es;
    int device;
    long long global[3];
    void* f[4];             // bound as u, v, u_temp, v_temp (swap exchanges 0<->2, 1<->3)
    CUtensorMap map[4];     // tensor map of f[i] for the current TMA tile height
    int map_ty;             // tile height the maps were encoded for (0 = none)
    cudaStream_t s_compute, s_comm;
    cudaStream_t s_compute2;  // pipelined scheme: chunks alternate between two update streams so that the
                              // next chunk fills the SMs the previous one leaves idle in its tail
    cudaEvent_t ev_join;
    // asynchronous output (gsb_get_fields_async): device snapshot + D2H on its own stream
    cudaStream_t s_copy;
    cudaEvent_t ev_snap, ev_copied;
    void* stage[2];          // two contiguous staging buffers of one ghosted field each (allocated on first use)
    bool stage_failed;       // the allocation was tried and did not fit
    bool copy_in_flight;
    cudaEvent_t ev_compute, ev_halo, ev_packed, ev_chunk;
    cudaEvent_t ev_piece[GSB_MAX_CHUNKS];   // pipelined halo: piece k of the previous step has landed
    int halo_mode, halo_chunks, fuse_x;
    bool ghost_valid;       // the ghosts of the current fields already hold the neighbours' current values
    long long step;
    StencilCfg kcfg;
    // halo
    bool active[6];
    int n_active;
    void* send[6];
    void* recv[2][6];       // double-buffered by step parity
    int transport;
    gsb_handle* peer[6];
    ncclComm_t comm;
    // scratch for gsb_field_stats
    double* stats_partials;
    double* stats_out;
    // timers
    int timers_on;
    cudaEvent_t tev[6];     // comm begin/end, compute begin, interior end, compute end, (spare)
    bool timers_valid;
    // per-piece halo timers: [piece][0..3] = before pack, after pack, after the transport, after unpack
    cudaEvent_t pev[GSB_MAX_CHUNKS + 1][4];
    long long piece_bytes[GSB_MAX_CHUNKS + 1];   // bytes this rank sent with the piece
    int pieces_timed;       // pieces recorded by the last step (0: none)
    int piece_cursor;       // slot the next enqueue_pack_send records into
    unsigned* pdf_dev;      // gsb_pdf: device histogram, grown on demand
    size_t pdf_cap;
    long long launches;     // kernels launched by this handle (bench.py's gpu_launches)
};

static inline int opp(int d) { return d ^ 1; }

// timing experiments only ($GSB_DEBUG_SKIP bit 0: no pack kernel, bit 1: no NCCL, bit 2: no unpack kernel):
// results are wrong when set
static int debug_skip() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("GSB_DEBUG_SKIP"); v = e ? atoi(e) : 0; }
    return v;
}

static Params make_params(const gsb_handle* h) {
    const gsb_config& c = h->cfg;
    Params p;
    memset(&p, 0, sizeof p);
    if (c.dtype == GSB_F32) {
        const float Du = (float)c.Du, Dv = (float)c.Dv, F = (float)c.F, K = (float)c.k, dt = (float)c.dt,
                    nz = (float)c.noise;                     // convert(T, settings.x)  Simulation_CPU.jl:82-87
        const float FK = F + K;                              // Float32 add, :105
        p.Du = Du; p.Dv = Dv; p.F = F; p.dt = dt; p.noise = nz; p.FK = FK;
        p.fDu = Du; p.fDv = Dv; p.fF = F; p.fFK = FK; p.fdt = dt; p.fnoise = nz;
        p.dt_is_one = dt == 1.0f;
    } else {
        p.Du = c.Du; p.Dv = c.Dv; p.F = c.F; p.dt = c.dt; p.noise = c.noise; p.FK = c.F + c.k;
        p.fDu = (float)c.Du; p.fDv = (float)c.Dv; p.fF = (float)c.F; p.fFK = (float)(c.F + c.k);
        p.fdt = (float)c.dt; p.fnoise = (float)c.noise;
        p.dt_is_one = c.dt == 1.0;
    }
    p.noise_s = ldexp(p.noise, -31);
    p.noise_fast = p.noise == 0.0 || (std::isnormal(p.noise_s) && ldexp(p.noise_s, 31) == p.noise);
    p.ref_mode = !p.noise_fast ? 2 : (p.dt_is_one ? 0 : 1);
    p.fDu6 = p.fDu / 6.0f;
    p.fDv6 = p.fDv / 6.0f;
    p.fnoise_s = ldexpf(p.fnoise, -23);
    p.seed_lo = (unsigned)(c.seed & 0xffffffffu);
    p.seed_hi = (unsigned)(c.seed >> 32);
    for (unsigned r = 0; r < 10; ++r) {
        p.pk[r][0] = p.seed_lo + r * 0x9E3779B9u;
        p.pk[r][1] = p.seed_hi + r * 0xBB67AE85u;
    }
    p.step = (unsigned)h->step;
    p.gx0 = (int)c.offsets[0]; p.gy0 = (int)c.offsets[1]; p.gz0 = (int)c.offsets[2];
    return p;
}

static int ensure_maps(gsb_handle* h) {
    if (h->kcfg.kind == KERNEL_L1) return 0;
    if (h->map_ty == h->kcfg.ty) return 0;
    for (int i = 0; i < 4; ++i) {
        const char* err = nullptr;
        if (make_field_tensor_map(&h->map[i], h->cfg.dtype, h->f[i], h->L, h->kcfg.ty, &err))
            return fail("tensor map: %s", err ? err : "?");
    }
    h->map_ty = h->kcfg.ty;
    return 0;
}

static void swap_bindings(gsb_handle* h) {
    std::swap(h->f[0], h->f[2]);
    std::swap(h->f[1], h->f[3]);
    std::swap(h->map[0], h->map[2]);
    std::swap(h->map[1], h->map[3]);
    h->step += 1;
}

static bool kernel_shape_ok(const StencilCfg& k) {
    return k.kind != KERNEL_L1
               ? ((k.ty == 4 && (k.stages == 4 || k.stages == 8)) || (k.ty == 8 && (k.stages == 3 || k.stages == 4 || k.stages == 6)) ||
                  (k.ty == 16 && (k.stages == 3 || k.stages == 4)))
               : (k.ty == 4 || k.ty == 8);
}

// Measured best shapes on B200 (profiles/r1_tuning_raw.log, profiles/r2_tuning_raw.log): occupancy class, rows per CTA, ring depth,
// planes per CTA; small grids get shorter z chunks so that every SM still holds several CTAs.
static void default_kernel(gsb_handle* h) {
    StencilCfg& k = h->kcfg;
    const bool nz = k.noise != 0;
    if (k.dtype == DT_F64) {
        k.kind = KERNEL_TMA_LO;
        if (nz) { k.ty = 4; k.stages = 4; k.zchunk = 32; } else { k.ty = 8; k.stages = 6; k.zchunk = 32; }
    } else if (k.arith == ARITH_NATIVE) {
        k.kind = KERNEL_TMA_HI; k.ty = 8;
        // with noise: 4 ring slots win the 20-step measurement by 5 % (2.89 against 3.05 ms at L = 1024) and lose
        // 1 % over 300 steps (3.04 against 3.01 ms) since the consumer loop was trimmed (profiles/r2_tuning_raw.log)
        if (nz) { k.stages = 4; k.zchunk = 32; } else { k.stages = 4; k.zchunk = 16; }
    } else {
        // Float32 reference arithmetic (FP64-pipe and conversion bound); re-measured after the consumer loop
        // lost its prologue test and most of its spills: the higher occupancy class wins again
        k.kind = KERNEL_TMA_HI; k.ty = 8; k.stages = 4; k.zchunk = 64;
    }
    // small grids: more, smaller CTAs (measured at L = 256: 4 rows per CTA beats 8 by 6 %)
    if ((long long)h->L.sx * h->L.sy * h->L.sz <= 300ll * 300 * 300 && (k.dtype == DT_F64 || k.arith == ARITH_NATIVE)) {
        k.ty = 4; k.stages = 4; k.zchunk = 16;
    }
    const int n = (int)(16 / h->es);
    const long long cells = (long long)h->L.sx * h->L.sy * h->L.sz;
    if (cells <= (k.dtype == DT_F64 ? 100ll * 100 * 100 : 140ll * 140 * 140)) {
        // L2-resident grids are latency-bound: the kernel without the TMA ring, 4 rows per CTA and very
        // short z chunks so that a few hundred CTAs exist (measured: L = 64 5.1 us per step against 13.3,
        // L = 128 10.3 against 14.4)
        k.kind = KERNEL_L1; k.ty = 4; k.stages = 4; k.zchunk = 32;   // (stages: unused by this kernel)
        const long long t4 = (long long)((h->L.sx + 32 * n - 1) / (32 * n)) * ((h->L.sy + 3) / 4);
        while (k.zchunk > 2 && t4 * ((h->L.sz + k.zchunk - 1) / k.zchunk) < 512) k.zchunk /= 2;
        return;
    }
    const long long tiles = (long long)((h->L.sx + 32 * n - 1) / (32 * n)) * ((h->L.sy + k.ty - 1) / k.ty);
    while (k.zchunk > 8 && tiles * ((h->L.sz + k.zchunk - 1) / k.zchunk) < 148 * 8) k.zchunk /= 2;
}

static int parse_kernel_env(StencilCfg* k) {
    const char* e = getenv("GSB_KERNEL");
    if (!e || !*e) return 0;
    char kind[16] = {0};
    int ty = 0, st = 0, zc = 0;
    const int n = sscanf(e, "%15[^:]:%d:%d:%d", kind, &ty, &st, &zc);
    if (n < 1) return fail("GSB_KERNEL: expected kind:ty:stages:zchunk, got '%s'", e);
    if (!strcmp(kind, "tma")) k->kind = KERNEL_TMA;
    else if (!strcmp(kind, "tmalo")) k->kind = KERNEL_TMA_LO;
    else if (!strcmp(kind, "tmahi")) k->kind = KERNEL_TMA_HI;
    else if (!strcmp(kind, "l1")) k->kind = KERNEL_L1;
    else return fail("GSB_KERNEL: unknown kind '%s' (tma|l1)", kind);
    if (n >= 2 && ty > 0) k->ty = ty;
    if (n >= 3 && st > 0) k->stages = st;
    if (n >= 4 && zc > 0) k->zchunk = zc;
    if (!kernel_shape_ok(*k)) return fail("GSB_KERNEL='%s': shape is not instantiated", e);
    return 0;
}

// ------------------------------------------------------------------ lifetime
extern "C" int gsb_version(void) { return GSB_VERSION; }
extern "C" const char* gsb_last_error(void) { return g_err.c_str(); }

extern "C" int gsb_dims_create(int n, int32_t dims[3]) {
    if (n < 1) return fail("gsb_dims_create: nranks must be >= 1");
    dims[0] = dims[1] = dims[2] = 1;
    int f[64], nf = 0, m = n;
    for (int p = 2; p * p <= m; ++p) while (m % p == 0) { f[nf++] = p; m /= p; }
    if (m > 1) f[nf++] = m;
    for (int i = nf - 1; i >= 0; --i) {
        int s = 0;
        for (int d = 1; d < 3; ++d) if (dims[d] < dims[s]) s = d;
        dims[s] *= f[i];
    }
    for (int a = 0; a < 3; ++a) for (int b = a + 1; b < 3; ++b) if (dims[b] > dims[a]) std::swap(dims[a], dims[b]);
    return 0;
}

extern "C" int gsb_fill_domain(gsb_config* c, int64_t L, int nranks, int rank) {
    if (!c) return fail("gsb_fill_domain: null config");
    if (rank < 0 || rank >= nranks) return fail("gsb_fill_domain: rank %d outside [0,%d)", rank, nranks);
    if (gsb_dims_create(nranks, c->dims)) return 1;
    c->L = L; c->rank = rank; c->nranks = nranks;
    const int dy = c->dims[1], dz = c->dims[2];
    c->coords[0] = rank / (dy * dz); c->coords[1] = (rank / dz) % dy; c->coords[2] = rank % dz;  // row-major
    for (int d = 0; d < 3; ++d) {
        if (L % c->dims[d] != 0)
            return fail("L = %lld does not divide by dims[%d] = %d (reference: InexactError, communication.jl:73)",
                        (long long)L, d, c->dims[d]);
        c->global[d] = L;
        c->sizes[d] = L / c->dims[d];
        c->offsets[d] = c->sizes[d] * c->coords[d];
    }
    auto rk = [&](int cx, int cy, int cz) -> int {
        if (cx < 0 || cy < 0 || cz < 0 || cx >= c->dims[0] || cy >= c->dims[1] || cz >= c->dims[2]) return -1;
        return (cx * dy + cy) * dz + cz;
    };
    const int cx = c->coords[0], cy = c->coords[1], cz = c->coords[2];
    c->neighbors[GSB_WEST] = rk(cx - 1, cy, cz); c->neighbors[GSB_EAST] = rk(cx + 1, cy, cz);
    c->neighbors[GSB_DOWN] = rk(cx, cy - 1, cz); c->neighbors[GSB_UP] = rk(cx, cy + 1, cz);
    c->neighbors[GSB_SOUTH] = rk(cx, cy, cz - 1); c->neighbors[GSB_NORTH] = rk(cx, cy, cz + 1);
    return 0;
}

extern "C" int gsb_destroy(gsb_t* h) {
    if (!h) return 0;
    DeviceGuard guard_(h->device);
    cudaDeviceSynchronize();
    if (h->comm && g_nccl.lib) g_nccl.CommDestroy(h->comm);
    for (int i = 0; i < 4; ++i) if (h->f[i]) cudaFree(h->f[i]);
    for (int d = 0; d < 6; ++d) {
        if (h->send[d]) cudaFree(h->send[d]);
        for (int q = 0; q < 2; ++q) if (h->recv[q][d]) cudaFree(h->recv[q][d]);
    }
    if (h->stats_partials) cudaFree(h->stats_partials);
    if (h->stats_out) cudaFree(h->stats_out);
    if (h->s_compute) cudaStreamDestroy(h->s_compute);
    if (h->s_comm) cudaStreamDestroy(h->s_comm);
    if (h->s_compute2) cudaStreamDestroy(h->s_compute2);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->s_copy) cudaStreamDestroy(h->s_copy);
    if (h->ev_snap) cudaEventDestroy(h->ev_snap);
    if (h->ev_copied) cudaEventDestroy(h->ev_copied);
    for (int i = 0; i < 2; ++i) if (h->stage[i]) cudaFree(h->stage[i]);
    if (h->ev_compute) cudaEventDestroy(h->ev_compute);
    if (h->ev_halo) cudaEventDestroy(h->ev_halo);
    if (h->ev_packed) cudaEventDestroy(h->ev_packed);
    if (h->ev_chunk) cudaEventDestroy(h->ev_chunk);
    for (int k = 0; k < GSB_MAX_CHUNKS; ++k) if (h->ev_piece[k]) cudaEventDestroy(h->ev_piece[k]);
    for (int i = 0; i < 6; ++i) if (h->tev[i]) cudaEventDestroy(h->tev[i]);
    for (int k = 0; k <= GSB_MAX_CHUNKS; ++k)
        for (int q = 0; q < 4; ++q) if (h->pev[k][q]) cudaEventDestroy(h->pev[k][q]);
    if (h->pdf_dev) cudaFree(h->pdf_dev);
    delete h;
    return 0;
}

static int create_impl(gsb_handle* h, const gsb_config* cfg) {
    h->cfg = *cfg;
    const gsb_config& c = h->cfg;
    if (c.dtype != GSB_F32 && c.dtype != GSB_F64) return fail("dtype must be GSB_F32 or GSB_F64");
    if (c.arith != GSB_ARITH_REFERENCE && c.arith != GSB_ARITH_NATIVE) return fail("unknown arith %d", c.arith);
    if (c.arith == GSB_ARITH_NATIVE && c.dtype != GSB_F32) return fail("GSB_ARITH_NATIVE exists for Float32 only");
    for (int d = 0; d < 3; ++d) {
        if (c.sizes[d] < 1 || c.sizes[d] > (1 << 30)) return fail("sizes[%d] = %lld out of range", d, (long long)c.sizes[d]);
        h->global[d] = c.global[d] > 0 ? c.global[d] : c.L;
        if (h->global[d] < 1) return fail("L / global extent must be >= 1");
        if (h->global[d] % 2 != 0)
            return fail("global extent %lld is odd (reference: Int64(L/2 - 6) raises InexactError, Simulation_CPU.jl:34)",
                        (long long)h->global[d]);
        if (c.offsets[d] < 0 || c.offsets[d] + c.sizes[d] > h->global[d])
            return fail("offsets[%d] + sizes[%d] exceeds the global extent", d, d);
        if (h->global[d] >= (1ll << 31)) return fail("global extent too large for the 32-bit cell coordinates of the noise counter");
    }
    h->es = c.dtype == GSB_F32 ? 4 : 8;
    h->device = c.device;
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, h->device));
    if (prop.major < 10) return fail("device %d is sm_%d%d; this library holds sm_100a code only", h->device, prop.major, prop.minor);
    h->L = make_layout((int)c.sizes[0], (int)c.sizes[1], (int)c.sizes[2], h->es);
    const size_t bytes = (size_t)h->L.elems() * h->es;
    for (int i = 0; i < 4; ++i) CK(cudaMalloc(&h->f[i], bytes));
    // the halo path (pack, NCCL, unpack) must not queue behind the thousands of CTAs of the interior
    // update: highest priority for the communication stream, lowest for the update stream
    int prio_lo = 0, prio_hi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CK(cudaStreamCreateWithPriority(&h->s_compute, cudaStreamNonBlocking, prio_lo));
    CK(cudaStreamCreateWithPriority(&h->s_comm, cudaStreamNonBlocking, prio_hi));
    CK(cudaStreamCreateWithPriority(&h->s_compute2, cudaStreamNonBlocking, prio_lo));
    CK(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&h->s_copy, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&h->ev_snap, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->ev_copied, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->ev_compute, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->ev_halo, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->ev_packed, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->ev_chunk, cudaEventDisableTiming));
    for (int k = 0; k < GSB_MAX_CHUNKS; ++k) CK(cudaEventCreateWithFlags(&h->ev_piece[k], cudaEventDisableTiming));
    {
        const char* hm = getenv("GSB_HALO");          // pipeline (default) | split
        h->halo_mode = (hm && !strcmp(hm, "split")) ? HALO_SPLIT : HALO_PIPELINE;
        const char* hc = getenv("GSB_HALO_CHUNKS");
        h->halo_chunks = hc ? atoi(hc) : 0;
        const char* fx = getenv("GSB_FUSE_X");
        h->fuse_x = fx ? atoi(fx) : 1;
    }
    for (int i = 0; i < 6; ++i) CK(cudaEventCreate(&h->tev[i]));
    for (int k = 0; k <= GSB_MAX_CHUNKS; ++k)
        for (int q = 0; q < 4; ++q) CK(cudaEventCreate(&h->pev[k][q]));
    h->n_active = 0;
    for (int d = 0; d < 6; ++d) {
        h->active[d] = c.neighbors[d] >= 0;
        if (h->active[d]) {
            h->n_active++;
            const size_t fb = 2 * (size_t)face_elems(h->L, d) * h->es;
            CK(cudaMalloc(&h->send[d], fb));
            CK(cudaMalloc(&h->recv[0][d], fb));
            CK(cudaMalloc(&h->recv[1][d], fb));
        }
    }
    CK(cudaMalloc(&h->stats_partials, sizeof(double) * (size_t)stats_scratch_doubles()));
    CK(cudaMalloc(&h->stats_out, sizeof(double) * 5));
    h->kcfg.dtype = c.dtype; h->kcfg.arith = c.arith; h->kcfg.noise = c.noise != 0.0;
    default_kernel(h);
    if (parse_kernel_env(&h->kcfg)) return 1;
    return ensure_maps(h);
}

extern "C" int gsb_create(gsb_t** out, const gsb_config* cfg) {
    if (!out || !cfg) return fail("gsb_create: null argument");
    *out = nullptr;
    gsb_handle* h = new gsb_handle();
    memset((void*)h, 0, sizeof *h);
    if (create_impl(h, cfg)) {
        std::string keep = g_err;
        gsb_destroy(h);
        g_err = keep;
        return 1;
    }
    *out = h;
    return 0;
}

// ------------------------------------------------------------------ init / access
extern "C" int gsb_init_fields(gsb_t* h) {
    if (!h) return fail("null handle");
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    const long long off[3] = {h->cfg.offsets[0], h->cfg.offsets[1], h->cfg.offsets[2]};
    CK(cudaStreamSynchronize(h->s_comm));      // no halo piece may still be landing in the ghosts
    CK(cudaStreamSynchronize(h->s_compute));
    CK(launch_init_fields(h->cfg.dtype, h->f[0], h->f[1], h->f[2], h->f[3], h->L, h->global, off, h->s_compute));
    CK(cudaEventRecord(h->ev_compute, h->s_compute));
    h->step = 0;
    h->ghost_valid = false;
    return 0;
}

extern "C" int gsb_sync(gsb_t* h) {
    if (!h) return fail("null handle");
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    CK(cudaStreamSynchronize(h->s_comm));
    CK(cudaStreamSynchronize(h->s_compute));
    return 0;
}

// Two contiguous staging buffers of one GHOSTED field each, allocated on first use.  Host transfers go
// through them: one dense cudaMemcpyAsync at full PCIe rate plus a strided device-to-device copy (a few
// ms at L = 1024) instead of one strided host copy (measured: 41-44 GB/s strided against >50 dense).
// Returns false when the memory is not there (L = 2048 on one GPU): callers then copy strided.
static bool ensure_stage(gsb_handle* h) {
    if (h->stage[0] && h->stage[1]) return true;
    if (h->stage_failed) return false;
    const Layout& L = h->L;
    const size_t bytes = (size_t)(L.sx + 2) * (size_t)(L.sy + 2) * (size_t)(L.sz + 2) * h->es;
    for (int i = 0; i < 2; ++i)
        if (!h->stage[i] && cudaMalloc(&h->stage[i], bytes) != cudaSuccess) {
            cudaGetLastError();
            for (int q = 0; q < 2; ++q) if (h->stage[q]) { cudaFree(h->stage[q]); h->stage[q] = nullptr; }
            h->stage_failed = true;
            return false;
        }
    return true;
}

// strided copy between the padded device array and a dense array `other` (g = 1: with ghosts) that lives
// on the host (kind H2D / D2H) or on the device (D2D: the staging buffers)
static int copy_strided(gsb_handle* h, void* dev, void* other, int g, bool from_dev, cudaMemcpyKind kind, cudaStream_t st) {
    const Layout& L = h->L;
    cudaMemcpy3DParms p;
    memset(&p, 0, sizeof p);
    const size_t nx = (size_t)(L.sx + 2 * g), ny = (size_t)(L.sy + 2 * g), nz = (size_t)(L.sz + 2 * g);
    // no CUDA array takes part: every x quantity is in bytes
    cudaPitchedPtr dptr = make_cudaPitchedPtr(dev, (size_t)L.px * h->es, (size_t)L.px * h->es, (size_t)(L.sy + 2));
    cudaPitchedPtr optr = make_cudaPitchedPtr(other, nx * h->es, nx * h->es, ny);
    cudaPos dpos = make_cudaPos((size_t)(L.xoff - g) * h->es, (size_t)(1 - g), (size_t)(1 - g));
    cudaPos opos = make_cudaPos(0, 0, 0);
    p.extent = make_cudaExtent(nx * h->es, ny, nz);
    if (from_dev) { p.srcPtr = dptr; p.srcPos = dpos; p.dstPtr = optr; p.dstPos = opos; }
    else { p.srcPtr = optr; p.srcPos = opos; p.dstPtr = dptr; p.dstPos = dpos; }
    p.kind = kind;
    CK(cudaMemcpy3DAsync(&p, st));
    return 0;
}

// padded device array <-> dense host array, on the update stream; `slot` picks the staging buffer
static int copy_field(gsb_handle* h, void* dev, void* host, int g, bool to_host, int slot = 0) {
    const Layout& L = h->L;
    const size_t bytes = (size_t)(L.sx + 2 * g) * (size_t)(L.sy + 2 * g) * (size_t)(L.sz + 2 * g) * h->es;
    // small fields: one strided copy has the lower latency
    if (bytes < ((size_t)8 << 20) || !ensure_stage(h))
        return copy_strided(h, dev, host, g, to_host, to_host ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice, h->s_compute);
    if (h->copy_in_flight) {                                 // an asynchronous output still drains from the buffers
        CK(cudaEventSynchronize(h->ev_copied));
        h->copy_in_flight = false;
    }
    void* stg = h->stage[slot & 1];
    if (to_host) {
        if (copy_strided(h, dev, stg, g, true, cudaMemcpyDeviceToDevice, h->s_compute)) return 1;
        CK(cudaMemcpyAsync(host, stg, bytes, cudaMemcpyDeviceToHost, h->s_compute));
    } else {
        CK(cudaMemcpyAsync(stg, host, bytes, cudaMemcpyHostToDevice, h->s_compute));
        if (copy_strided(h, dev, stg, g, false, cudaMemcpyDeviceToDevice, h->s_compute)) return 1;
    }
    return 0;
}

static int refresh_x_ghosts(gsb_handle* h);

extern "C" int gsb_get_fields(gsb_t* h, void* u_host, void* v_host) {
    if (!h || !u_host || !v_host) return fail("gsb_get_fields: null argument");
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    CK(cudaStreamSynchronize(h->s_comm));
    if (copy_field(h, h->f[0], u_host, 0, true, 0)) return 1;
    if (copy_field(h, h->f[1], v_host, 0, true, 1)) return 1;
    CK(cudaStreamSynchronize(h->s_compute));
    return 0;
}

// Output without stalling the time loop (IO.write_step! every plotgap steps, IO.jl:82-96): a ghost-free
// snapshot of u and v is taken on the update stream (device to device), then drained to the host on
// a copy stream while later steps run.
extern "C" int gsb_output_wait(gsb_t* h) {
    if (!h) return fail("null handle");
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    if (h->copy_in_flight) {
        CK(cudaEventSynchronize(h->ev_copied));
        h->copy_in_flight = false;
    }
    return 0;
}

extern "C" int gsb_get_fields_async(gsb_t* h, void* u_host, void* v_host) {
    if (!h || !u_host || !v_host) return fail("gsb_get_fields_async: null argument");
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    const Layout& L = h->L;
    const size_t nx = (size_t)L.sx, ny = (size_t)L.sy, nz = (size_t)L.sz;
    const size_t bytes = nx * ny * nz * h->es;
    if (!ensure_stage(h)) return gsb_get_fields(h, u_host, v_host);   // no room for a snapshot (L = 2048 on one GPU)
    if (gsb_output_wait(h)) return 1;               // the staging buffers are free again
    void* host[2] = {u_host, v_host};
    for (int i = 0; i < 2; ++i) {
        cudaMemcpy3DParms p;
        memset(&p, 0, sizeof p);
        p.srcPtr = make_cudaPitchedPtr(h->f[i], (size_t)L.px * h->es, (size_t)L.px * h->es, (size_t)(L.sy + 2));
        p.srcPos = make_cudaPos((size_t)L.xoff * h->es, 1, 1);
        p.dstPtr = make_cudaPitchedPtr(h->stage[i], nx * h->es, nx * h->es, ny);
        p.dstPos = make_cudaPos(0, 0, 0);
        p.extent = make_cudaExtent(nx * h->es, ny, nz);
        p.kind = cudaMemcpyDeviceToDevice;
        CK(cudaMemcpy3DAsync(&p, h->s_compute));
    }
    CK(cudaEventRecord(h->ev_snap, h->s_compute));
    CK(cudaStreamWaitEvent(h->s_copy, h->ev_snap, 0));
    for (int i = 0; i < 2; ++i) CK(cudaMemcpyAsync(host[i], h->stage[i], bytes, cudaMemcpyDeviceToHost, h->s_copy));
    CK(cudaEventRecord(h->ev_copied, h->s_copy));
    h->copy_in_flight = true;
    return 0;
}

extern "C" int gsb_get_ghosted(gsb_t* h, int which, void* host) {
    if (!h || !host) return fail("gsb_get_ghosted: null argument");
    if (which < 0 || which > 3) return fail("gsb_get_ghosted: which = %d", which);
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    CK(cudaStreamSynchronize(h->s_comm));
    if (refresh_x_ghosts(h)) return 1;
    if (copy_field(h, h->f[which], host, 1, true)) return 1;
    CK(cudaStreamSynchronize(h->s_compute));
    return 0;
}

extern "C" int gsb_set_ghosted(gsb_t* h, int which, const void* host) {
    if (!h || !host) return fail("gsb_set_ghosted: null argument");
    if (which < 0 || which > 3) return fail("gsb_set_ghosted: which = %d", which);
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    CK(cudaStreamSynchronize(h->s_comm));
    if (copy_field(h, h->f[which], const_cast<void*>(host), 1, false)) return 1;
    h->ghost_valid = false;
    CK(cudaEventRecord(h->ev_compute, h->s_compute));
    CK(cudaStreamSynchronize(h->s_compute));
    return 0;
}

extern "C" int gsb_alloc_host(void** out, size_t bytes) {
    if (!out) return fail("gsb_alloc_host: null argument");
    CK(cudaHostAlloc(out, bytes, cudaHostAllocDefault));
    return 0;
}
extern "C" int gsb_free_host(void* p) {
    if (p) CK(cudaFreeHost(p));
    return 0;
}

extern "C" int gsb_get_step(gsb_t* h, int64_t* step) {
    if (!h || !step) return fail("null argument");
    *step = h->step;
    return 0;
}
extern "C" int gsb_set_step(gsb_t* h, int64_t step) {
    if (!h) return fail("null handle");
    if (step < 0) return fail("gsb_set_step: step < 0");
    h->step = step;
    h->ghost_valid = false;   // receive buffers and message parity follow the step number: exchange again
    return 0;
}

extern "C" int gsb_field_stats(gsb_t* h, int which, double thresh, double out[5]) {
    if (!h || !out) return fail("null argument");
    if (which < 0 || which > 3) return fail("gsb_field_stats: which = %d", which);
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    CK(cudaStreamSynchronize(h->s_comm));
    CK(launch_field_stats(h->cfg.dtype, h->f[which], h->L, thresh, h->stats_partials, h->stats_out, h->s_compute));
    CK(cudaMemcpyAsync(out, h->stats_out, sizeof(double) * 5, cudaMemcpyDeviceToHost, h->s_compute));
    CK(cudaStreamSynchronize(h->s_compute));
    return 0;
}

// _compute_pdf of src/analysis/pdfcalc.jl:13-49 on the resident field
extern "C" int gsb_pdf(gsb_t* h, int which, int nbins, double vmin, double vmax, uint32_t* pdf_host, double* bins_host) {
    if (!h || !pdf_host) return fail("gsb_pdf: null argument");
    if (which < 0 || which > 3) return fail("gsb_pdf: which = %d", which);
    if (nbins < 1 || nbins > 8192) return fail("gsb_pdf: nbins = %d outside [1, 8192]", nbins);
    if (!(vmax >= vmin)) return fail("gsb_pdf: max < min");
    const size_t n = (size_t)nbins * (size_t)h->L.sz;
    if (bins_host) {
        // bins[i] = min + i * bin_width, evaluated in T like the reference
        if (h->cfg.dtype == GSB_F32) {
            const float lo = (float)vmin, w = ((float)vmax - lo) / (float)nbins;
            for (int i = 0; i < nbins; ++i) bins_host[i] = (double)(lo + (float)i * w);
        } else {
            const double w = (vmax - vmin) / (double)nbins;
            for (int i = 0; i < nbins; ++i) bins_host[i] = vmin + (double)i * w;
        }
    }
    // special case of the reference: one bin or an empty window -> every slice holds slice_size
    const double width = (vmax - vmin) / nbins;
    if (nbins == 1 || (vmax - vmin) < 1.0e-20 || width < 1.0e-20) {
        for (size_t i = 0; i < n; ++i) pdf_host[i] = (uint32_t)((size_t)h->L.sx * h->L.sy);
        return 0;
    }
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    CK(cudaStreamSynchronize(h->s_comm));
    if (h->pdf_cap < n) {                                   // histogram buffer kept by the handle, grown on demand
        if (h->pdf_dev) { CK(cudaFree(h->pdf_dev)); h->pdf_dev = nullptr; h->pdf_cap = 0; }
        CK(cudaMalloc((void**)&h->pdf_dev, n * sizeof(unsigned)));
        h->pdf_cap = n;
    }
    unsigned* dev = h->pdf_dev;
    cudaError_t e = launch_pdf(h->cfg.dtype, h->f[which], h->L, nbins, vmin, vmax, dev, h->s_compute);
    if (e == cudaSuccess) e = cudaMemcpyAsync(pdf_host, dev, n * sizeof(unsigned), cudaMemcpyDeviceToHost, h->s_compute);
    if (e != cudaSuccess) return fail("gsb_pdf: %s", cudaGetErrorString(e));
    CK(cudaStreamSynchronize(h->s_compute));
    return 0;
}

// ------------------------------------------------------------------ kernel selection, timers
extern "C" int gsb_set_kernel(gsb_t* h, int kind, int ty, int stages, int zchunk) {
    if (!h) return fail("null handle");
    StencilCfg k = h->kcfg;
    if (kind >= GSB_KERNEL_TMA && kind <= GSB_KERNEL_TMA_HI) k.kind = kind;
    else if (kind >= 0) return fail("unknown kernel kind %d", kind);
    if (ty > 0) k.ty = ty;
    if (stages > 0) k.stages = stages;
    if (zchunk > 0) k.zchunk = zchunk;
    if (!kernel_shape_ok(k)) return fail("kernel shape kind=%d ty=%d stages=%d is not instantiated", k.kind, k.ty, k.stages);
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    h->kcfg = k;
    h->ghost_valid = false;   // the kernel kind decides who carries the x faces
    return ensure_maps(h);
}
extern "C" int gsb_get_kernel(gsb_t* h, int* kind, int* ty, int* stages, int* zchunk) {
    if (!h) return fail("null handle");
    if (kind) *kind = h->kcfg.kind;
    if (ty) *ty = h->kcfg.ty;
    if (stages) *stages = h->kcfg.stages;
    if (zchunk) *zchunk = h->kcfg.zchunk;
    return 0;
}
extern "C" int gsb_timers_enable(gsb_t* h, int on) {
    if (!h) return fail("null handle");
    h->timers_on = on;
    h->timers_valid = false;
    return 0;
}
extern "C" int gsb_timers(gsb_t* h, double out_ms[4]) {
    if (!h || !out_ms) return fail("null argument");
    if (!h->timers_on || !h->timers_valid) return fail("gsb_timers: enable timers and run gsb_step first");
    if (gsb_sync(h)) return 1;
    float a = 0, b = 0, c = 0, d = 0;
    for (int k = 0; k < h->pieces_timed; ++k) {                 // pack -> transport -> unpack, every piece of the step
        float t = 0;
        CK(cudaEventElapsedTime(&t, h->pev[k][0], h->pev[k][3]));
        a += t;
    }
    CK(cudaEventElapsedTime(&b, h->tev[2], h->tev[3]));
    CK(cudaEventElapsedTime(&c, h->tev[5], h->tev[4]));   // shell kernels only, after the halo arrived
    CK(cudaEventElapsedTime(&d, h->tev[2], h->tev[4]));
    out_ms[0] = a; out_ms[1] = b; out_ms[2] = c; out_ms[3] = d;
    return 0;
}
extern "C" int gsb_halo_timers(gsb_t* h, int max_pieces, double* pack_ms, double* wire_ms, double* unpack_ms,
                               int64_t* bytes, int* npieces) {
    if (!h || !npieces) return fail("null argument");
    if (!h->timers_on) return fail("gsb_halo_timers: enable timers and run gsb_step first");
    if (gsb_sync(h)) return 1;
    const int n = h->pieces_timed < max_pieces ? h->pieces_timed : max_pieces;
    for (int k = 0; k < n; ++k) {
        float p = 0, w = 0, u = 0;
        CK(cudaEventElapsedTime(&p, h->pev[k][0], h->pev[k][1]));
        CK(cudaEventElapsedTime(&w, h->pev[k][1], h->pev[k][2]));
        CK(cudaEventElapsedTime(&u, h->pev[k][2], h->pev[k][3]));
        if (pack_ms) pack_ms[k] = p;
        if (wire_ms) wire_ms[k] = w;
        if (unpack_ms) unpack_ms[k] = u;
        if (bytes) bytes[k] = h->piece_bytes[k];
    }
    *npieces = n;
    return 0;
}
extern "C" int gsb_device_count(int* count) {
    if (!count) return fail("null argument");
    CK(cudaGetDeviceCount(count));
    return 0;
}
extern "C" int gsb_halo_bytes(gsb_t* h, int64_t* bytes) {
    if (!h || !bytes) return fail("null argument");
    long long b = 0;
    for (int d = 0; d < 6; ++d) if (h->active[d]) b += 2 * face_elems(h->L, d) * (long long)h->es;
    *bytes = b;
    return 0;
}
extern "C" int gsb_launch_count(gsb_t* h, int64_t* count) {
    if (!h || !count) return fail("null argument");
    *count = h->launches;
    return 0;
}
extern "C" int gsb_compute_stream(gsb_t* h, void** stream) {
    if (!h || !stream) return fail("null argument");
    *stream = (void*)h->s_compute;
    return 0;
}

// ------------------------------------------------------------------ communication set-up
extern "C" int gsb_get_unique_id(void* id128) {
    if (!id128) return fail("null argument");
    if (nccl_load()) return 1;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    NK(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return 0;
}

extern "C" int gsb_comm_init(gsb_t* h, const void* id128) {
    if (!h || !id128) return fail("null argument");
    if (h->transport != TR_NONE) return fail("communication already initialised");
    if (nccl_load()) return 1;
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    NK(g_nccl.CommInitRank(&h->comm, h->cfg.nranks, id, h->cfg.rank));
    h->transport = TR_NCCL;
    return 0;
}

extern "C" int gsb_comm_init_local(gsb_t** hs, int n) {
    if (!hs || n < 1) return fail("gsb_comm_init_local: bad arguments");
    for (int r = 0; r < n; ++r) {
        if (!hs[r]) return fail("gsb_comm_init_local: null handle %d", r);
        if (hs[r]->cfg.nranks != n || hs[r]->cfg.rank != r)
            return fail("gsb_comm_init_local: hs[%d] has rank %d of %d", r, hs[r]->cfg.rank, hs[r]->cfg.nranks);
        if (hs[r]->transport != TR_NONE) return fail("communication already initialised");
    }
    for (int r = 0; r < n; ++r) {
        gsb_handle* h = hs[r];
        for (int d = 0; d < 6; ++d) {
            h->peer[d] = nullptr;
            if (!h->active[d]) continue;
            const int q = h->cfg.neighbors[d];
            if (q < 0 || q >= n) return fail("rank %d: neighbour %d out of range", r, q);
            gsb_handle* p = hs[q];
            if (p->cfg.neighbors[opp(d)] != r) return fail("rank %d and %d disagree about being neighbours", r, q);
            if (face_elems(p->L, opp(d)) != face_elems(h->L, d)) return fail("rank %d and %d: face sizes differ", r, q);
            h->peer[d] = p;
            if (p->device != h->device) {
                int can = 0;
                CK(cudaDeviceCanAccessPeer(&can, h->device, p->device));
                if (!can) return fail("device %d cannot access device %d", h->device, p->device);
                DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
                cudaError_t e = cudaDeviceEnablePeerAccess(p->device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else CK(e);
            }
        }
    }
    for (int r = 0; r < n; ++r) hs[r]->transport = TR_LOCAL;
    return 0;
}

// ------------------------------------------------------------------ step orchestration
static Box full_box(const gsb_handle* h) { return Box{0, h->L.sx, 0, h->L.sy, 0, h->L.sz}; }

static int launch_box(gsb_handle* h, const Box& b, const Params& p, cudaStream_t st = nullptr,
                      const XFaceArgs* xf = nullptr) {
    CK(launch_stencil(h->kcfg, h->f[0], h->f[1], h->f[2], h->f[3], &h->map[0], &h->map[1], h->L, p, b,
                      st ? st : h->s_compute, xf));
    h->launches += 1;
    return 0;
}

// thickness of the shell (cells next to an exchanged face) per axis: one tile
static void shell_thickness(const gsb_handle* h, int t[3]) {
    t[0] = 32 * (int)(16 / h->es);
    t[1] = h->kcfg.ty;
    t[2] = 1;
}

// the part of the update that reads no exchanged ghost cell
static bool interior_box(const gsb_handle* h, Box* b) {
    int t[3];
    shell_thickness(h, t);
    const int s[3] = {h->L.sx, h->L.sy, h->L.sz};
    int lo[3], hi[3];
    for (int d = 0; d < 3; ++d) {
        lo[d] = h->active[2 * d] ? t[d] : 0;
        hi[d] = h->active[2 * d + 1] ? s[d] - t[d] : s[d];
        if (hi[d] <= lo[d]) return false;
    }
    *b = Box{lo[0], hi[0], lo[1], hi[1], lo[2], hi[2]};
    return true;
}

// The x faces ride inside the update kernel (XFaceArgs, gs_stencil.cuh) when the pipelined scheme runs
// the TMA kernel on a subdomain whose sx is a whole number of tiles and whose sy is a whole number of
// 16-byte vectors: the producer warps of the first / last x tile fetch the neighbour's column from the
// dense receive buffer and send the new boundary column to the dense message buffer with bulk copies.
// That removes the strided pack / unpack of 2 * sy * sz cells (a 128-byte DRAM line per 4-byte cell).
// $GSB_FUSE_X at gsb_create: 1 (default) one launch per chunk, every CTA runs the x-face instantiation (it no
// longer spills, and only the edge tiles act on it); 2 the x-face instantiation on the first / last x tile only,
// the plain kernel in between; 0 off (the strided kernels then carry the x faces as well).
static bool fused_x(const gsb_handle* h) {
    const int tile = 32 * (int)(16 / h->es);
    // in-process ranks on DIFFERENT devices keep the strided pack kernel for the x faces: its plain stores into
    // the neighbour GPU's receive buffer are the tested peer path, a bulk store from shared memory to a peer is not
    for (int d = GSB_WEST; d <= GSB_EAST; ++d)
        if (h->transport == TR_LOCAL && h->active[d] && h->peer[d] && h->peer[d]->device != h->device) return false;
    return h->fuse_x != 0 && h->halo_mode == HALO_PIPELINE && h->kcfg.kind != KERNEL_L1 &&
           (h->active[GSB_WEST] || h->active[GSB_EAST]) && h->L.sx % tile == 0 && h->L.sy % (int)(16 / h->es) == 0;
}

// A piece of the halo: which faces travel, and for the x / y faces which z range.
struct Piece {
    int z0, z1;
    bool send[6], recv[6];
};

static Piece whole_piece(const gsb_handle* h) {
    Piece p;
    p.z0 = 0; p.z1 = h->L.sz;
    for (int d = 0; d < 6; ++d) p.send[d] = p.recv[d] = h->active[d];
    return p;
}

// byte offset of a piece inside the message buffer of face d ([u piece | v piece], pieces in z order)
static size_t piece_offset(const gsb_handle* h, int d, int z0) {
    if ((d >> 1) == 2) return 0;
    const long long row = (d >> 1) == 0 ? h->L.sy : h->L.sx;
    return (size_t)(2 * row * z0) * h->es;
}

// pack -> (transport): first half of an exchange, on the communication stream.
//   src = 0: the fields as bound now (u, v); src = 1: the fields being written (u_temp, v_temp)
//   par: which of the two receive buffers this exchange uses
static int enqueue_pack_send(gsb_handle* h, const Piece& pc, int src, int par, cudaEvent_t after, bool xfused = false) {
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    CK(cudaStreamWaitEvent(h->s_comm, after, 0));               // the cells to send are final
    const int slot = h->piece_cursor <= GSB_MAX_CHUNKS ? h->piece_cursor : GSB_MAX_CHUNKS;
    if (h->timers_on) {
        CK(cudaEventRecord(h->pev[slot][0], h->s_comm));
        long long b = 0;
        for (int d = 0; d < 6; ++d) if (pc.send[d]) b += 2 * face_piece_elems(h->L, d, pc.z0, pc.z1) * (long long)h->es;
        h->piece_bytes[slot] = b;
    }
    FaceSet fs;
    fs.n = 0;
    for (int d = 0; d < 6; ++d) {
        if (!pc.send[d]) continue;
        if (d < 2 && src == 1 && xfused) continue;              // the update kernel wrote this message itself
        char* base = (char*)(h->transport == TR_LOCAL ? h->peer[d]->recv[par][opp(d)] : h->send[d]);
        fs.dir[fs.n] = d;
        fs.buf[fs.n] = base + piece_offset(h, d, pc.z0);
        fs.z0[fs.n] = pc.z0; fs.z1[fs.n] = pc.z1;
        fs.n++;
    }
    if (fs.n && !(debug_skip() & 1)) {
        CK(launch_pack_faces(h->cfg.dtype, h->f[2 * src], h->f[2 * src + 1], h->L, fs, h->s_comm));
        h->launches += 1;
    }
    if (h->timers_on) CK(cudaEventRecord(h->pev[slot][1], h->s_comm));
    if (h->transport == TR_NCCL && !(debug_skip() & 2)) {
        const ncclDataType_t ty = h->cfg.dtype == GSB_F32 ? ncclFloat : ncclDouble;
        NK(g_nccl.GroupStart());
        for (int d = 0; d < 6; ++d) {
            const size_t cnt = 2 * (size_t)face_piece_elems(h->L, d, pc.z0, pc.z1);
            const size_t off = piece_offset(h, d, pc.z0);
            if (pc.send[d]) NK(g_nccl.Send((char*)h->send[d] + off, cnt, ty, h->cfg.neighbors[d], h->comm, h->s_comm));
            if (pc.recv[d]) NK(g_nccl.Recv((char*)h->recv[par][d] + off, cnt, ty, h->cfg.neighbors[d], h->comm, h->s_comm));
        }
        NK(g_nccl.GroupEnd());
    }
    if (h->timers_on) CK(cudaEventRecord(h->pev[slot][2], h->s_comm));
    CK(cudaEventRecord(h->ev_packed, h->s_comm));
    return 0;
}

// (wait for the neighbours) -> unpack into the ghosts of the fields `dst` (0: u, v; 1: u_temp, v_temp)
static int enqueue_unpack(gsb_handle* h, const Piece& pc, int dst, int par, cudaEvent_t done, bool xfused = false) {
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    if (h->transport == TR_LOCAL)
        for (int d = 0; d < 6; ++d)
            if (pc.recv[d]) CK(cudaStreamWaitEvent(h->s_comm, h->peer[d]->ev_packed, 0));
    FaceSet fs;
    fs.n = 0;
    for (int d = 0; d < 6; ++d) {
        if (!pc.recv[d]) continue;
        if (d < 2 && xfused) continue;                          // the update kernel reads the receive buffer
        fs.dir[fs.n] = d;
        fs.buf[fs.n] = (char*)h->recv[par][d] + piece_offset(h, d, pc.z0);
        fs.z0[fs.n] = pc.z0; fs.z1[fs.n] = pc.z1;
        fs.n++;
    }
    if (fs.n && !(debug_skip() & 4)) {
        CK(launch_unpack_faces(h->cfg.dtype, h->f[2 * dst], h->f[2 * dst + 1], h->L, fs, h->s_comm));
        h->launches += 1;
    }
    if (h->timers_on) {
        const int slot = h->piece_cursor <= GSB_MAX_CHUNKS ? h->piece_cursor : GSB_MAX_CHUNKS;
        CK(cudaEventRecord(h->pev[slot][3], h->s_comm));
        if (h->piece_cursor < GSB_MAX_CHUNKS) h->piece_cursor += 1;
        h->pieces_timed = h->piece_cursor;
    }
    CK(cudaEventRecord(done, h->s_comm));
    return 0;
}

static int check_transport(gsb_handle* h) {
    if (h->n_active && h->transport == TR_NONE)
        return fail("rank %d has neighbours but no transport: call gsb_comm_init or gsb_comm_init_local", h->cfg.rank);
    return 0;
}

// ---- scheme A ("split"): exchange at the start of the step, interior update meanwhile, shell after
static int enqueue_compute_overlapped(gsb_handle* h) {
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    if (ensure_maps(h)) return 1;
    const Params p = make_params(h);
    if (h->timers_on) CK(cudaEventRecord(h->tev[2], h->s_compute));
    Box in;
    const bool split = h->n_active > 0 && interior_box(h, &in);
    if (!split) {
        if (h->n_active) CK(cudaStreamWaitEvent(h->s_compute, h->ev_halo, 0));
        if (h->timers_on) { CK(cudaEventRecord(h->tev[3], h->s_compute)); CK(cudaEventRecord(h->tev[5], h->s_compute)); }
        if (launch_box(h, full_box(h), p)) return 1;
    } else {
        if (launch_box(h, in, p)) return 1;
        if (h->timers_on) CK(cudaEventRecord(h->tev[3], h->s_compute));
        CK(cudaStreamWaitEvent(h->s_compute, h->ev_halo, 0));
        if (h->timers_on) CK(cudaEventRecord(h->tev[5], h->s_compute));
        const int sx = h->L.sx, sy = h->L.sy, sz = h->L.sz;
        // disjoint slabs: z slabs span everything, y slabs the interior z range, x slabs the interior y,z range
        if (in.z0 > 0) if (launch_box(h, Box{0, sx, 0, sy, 0, in.z0}, p)) return 1;
        if (in.z1 < sz) if (launch_box(h, Box{0, sx, 0, sy, in.z1, sz}, p)) return 1;
        if (in.y0 > 0) if (launch_box(h, Box{0, sx, 0, in.y0, in.z0, in.z1}, p)) return 1;
        if (in.y1 < sy) if (launch_box(h, Box{0, sx, in.y1, sy, in.z0, in.z1}, p)) return 1;
        if (in.x0 > 0) if (launch_box(h, Box{0, in.x0, in.y0, in.y1, in.z0, in.z1}, p)) return 1;
        if (in.x1 < sx) if (launch_box(h, Box{in.x1, sx, in.y0, in.y1, in.z0, in.z1}, p)) return 1;
    }
    if (h->timers_on) { CK(cudaEventRecord(h->tev[4], h->s_compute)); h->timers_valid = true; }
    CK(cudaEventRecord(h->ev_compute, h->s_compute));
    return 0;
}

// exchange of the current fields, whole faces (exchange! as the reference orders it: before calculate!)
static int enqueue_exchange_now(gsb_handle* h, bool local_second_half) {
    const Piece w = whole_piece(h);
    const int par = (int)(h->step & 1);
    if (!local_second_half) return enqueue_pack_send(h, w, 0, par, h->ev_compute);
    return enqueue_unpack(h, w, 0, par, h->ev_halo);
}

// ---- scheme B ("pipeline", default): the update runs as C launches over z chunks of FULL rows; as
// soon as a chunk is written, its new boundary values travel to the neighbours and land in the
// ghosts of the fields being written -- i.e. the exchange! of the NEXT step, one step ahead.  A
// chunk of the next step waits only for its own piece, which by then is most of a step old.
static int pipeline_chunks(const gsb_handle* h) {
    if (!h->n_active) return 1;
    const int sz = h->L.sz;
    int c = h->halo_chunks > 0 ? h->halo_chunks : (sz >= 192 ? 4 : sz >= 32 ? 2 : 1);
    if (c > GSB_MAX_CHUNKS) c = GSB_MAX_CHUNKS;
    if (c > sz) c = sz;
    return c < 1 ? 1 : c;
}

// chunk visiting order: the two chunks that own a z face first, so their planes leave early
static void chunk_order(int C, int order[GSB_MAX_CHUNKS]) {
    int n = 0;
    order[n++] = C - 1;
    if (C > 1) order[n++] = 0;
    for (int k = 1; k < C - 1; ++k) order[n++] = k;
}

static Piece chunk_piece(const gsb_handle* h, int k, int C) {
    Piece p;
    p.z0 = (int)((long long)h->L.sz * k / C);
    p.z1 = (int)((long long)h->L.sz * (k + 1) / C);
    for (int d = 0; d < 4; ++d) p.send[d] = p.recv[d] = h->active[d];
    // my lowest plane goes south with chunk 0 and the south neighbour's answer (its highest plane)
    // comes with ITS last chunk; symmetrically for north
    p.send[GSB_SOUTH] = h->active[GSB_SOUTH] && k == 0;
    p.recv[GSB_NORTH] = h->active[GSB_NORTH] && k == 0;
    p.send[GSB_NORTH] = h->active[GSB_NORTH] && k == C - 1;
    p.recv[GSB_SOUTH] = h->active[GSB_SOUTH] && k == C - 1;
    return p;
}

// x-face arguments of chunk k for the update kernel (fused scheme); `step_par` = parity of the step
// whose ghosts are READ, the new values go to the other parity
static XFaceArgs chunk_xface(const gsb_handle* h, int k, int C) {
    XFaceArgs xf;
    memset(&xf, 0, sizeof xf);
    const Piece pc = chunk_piece(h, k, C);
    xf.zbase = pc.z0;
    xf.nz = pc.z1 - pc.z0;
    const int rd = (int)(h->step & 1), wr = (int)((h->step + 1) & 1);
    auto at = [&](void* base, int d) { return (void*)((char*)base + piece_offset(h, d, pc.z0)); };
    if (h->active[GSB_WEST]) {
        xf.ghost_lo = at(h->recv[rd][GSB_WEST], GSB_WEST);
        xf.send_lo = at(h->transport == TR_LOCAL ? h->peer[GSB_WEST]->recv[wr][GSB_EAST] : h->send[GSB_WEST], GSB_WEST);
    }
    if (h->active[GSB_EAST]) {
        xf.ghost_hi = at(h->recv[rd][GSB_EAST], GSB_EAST);
        xf.send_hi = at(h->transport == TR_LOCAL ? h->peer[GSB_EAST]->recv[wr][GSB_WEST] : h->send[GSB_EAST], GSB_EAST);
    }
    return xf;
}

// which piece event chunk k must see before it may read its ghosts (events complete in visiting order)
static int chunk_dependency(int k, int C) { return (k == 0 || k == C - 1) ? 0 : k; }

static int pipeline_compute_chunk(gsb_handle* h, int k, int C, const Params& p, int i) {
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    const bool first = i == 0, last = i == C - 1;
    cudaStream_t st = (i & 1) ? h->s_compute2 : h->s_compute;   // chunks are independent of one another
    if (first) {
        if (ensure_maps(h)) return 1;
        if (h->timers_on) CK(cudaEventRecord(h->tev[2], h->s_compute));
    }
    if (i == 1) CK(cudaStreamWaitEvent(h->s_compute2, h->ev_compute, 0));   // previous step is complete
    if (h->n_active) CK(cudaStreamWaitEvent(st, h->ev_piece[chunk_dependency(k, C)], 0));
    const Piece pc = chunk_piece(h, k, C);
    const int sx = h->L.sx, sy = h->L.sy;
    if (fused_x(h)) {
        const XFaceArgs xf = chunk_xface(h, k, C);
        const int tile = 32 * (int)(16 / h->es);
        if (h->fuse_x == 1 || sx <= tile) {
            // one launch, every CTA runs the x-face instantiation (only the edge tiles act on it)
            if (launch_box(h, Box{0, sx, 0, sy, pc.z0, pc.z1}, p, st, &xf)) return 1;
        } else {
            // the x-face instantiation on the first / last x tile only, the plain kernel in between (was the
            // better choice while that instantiation spilled at the register cap of the noisy variants)
            int lo = 0, hi = sx;
            if (xf.send_lo) {
                XFaceArgs a = xf; a.send_hi = nullptr; a.ghost_hi = nullptr;
                if (launch_box(h, Box{0, tile, 0, sy, pc.z0, pc.z1}, p, st, &a)) return 1;
                lo = tile;
            }
            if (xf.send_hi) {
                XFaceArgs a = xf; a.send_lo = nullptr; a.ghost_lo = nullptr;
                if (launch_box(h, Box{sx - tile, sx, 0, sy, pc.z0, pc.z1}, p, st, &a)) return 1;
                hi = sx - tile;
            }
            if (hi > lo) if (launch_box(h, Box{lo, hi, 0, sy, pc.z0, pc.z1}, p, st)) return 1;
        }
    } else if (launch_box(h, Box{0, sx, 0, sy, pc.z0, pc.z1}, p, st)) return 1;
    CK(cudaEventRecord(h->ev_chunk, st));
    if (last) {
        if (C > 1) {                                             // join the second stream into the first
            CK(cudaEventRecord(h->ev_join, h->s_compute2));
            CK(cudaStreamWaitEvent(h->s_compute, h->ev_join, 0));
        }
        if (h->timers_on) {
            CK(cudaEventRecord(h->tev[3], h->s_compute)); CK(cudaEventRecord(h->tev[5], h->s_compute));
            CK(cudaEventRecord(h->tev[4], h->s_compute));
            h->timers_valid = true;
        }
        CK(cudaEventRecord(h->ev_compute, h->s_compute));
    }
    return 0;
}

// bring the ghosts of the current fields up to date once (first step after init / set / piecewise use):
// the same pieces, in the same order and message layout as the steady state, but packed from the
// CURRENT fields
static int pipeline_prime(gsb_handle** hs, int n) {
    bool need = false;
    for (int r = 0; r < n; ++r) need = need || (hs[r]->n_active && !hs[r]->ghost_valid);
    if (!need) return 0;
    const int C = pipeline_chunks(hs[0]);
    int order[GSB_MAX_CHUNKS];
    chunk_order(C, order);
    for (int i = 0; i < C; ++i) {
        const int k = order[i];
        for (int r = 0; r < n; ++r) {
            gsb_handle* h = hs[r];
            if (h->n_active) if (enqueue_pack_send(h, chunk_piece(h, k, C), 0, (int)(h->step & 1), h->ev_compute)) return 1;
        }
        for (int r = 0; r < n; ++r) {
            gsb_handle* h = hs[r];
            if (h->n_active)
                if (enqueue_unpack(h, chunk_piece(h, k, C), 0, (int)(h->step & 1), h->ev_piece[k], fused_x(h))) return 1;
        }
    }
    for (int r = 0; r < n; ++r) if (hs[r]->n_active) {
        // unused piece events (k >= C) must not hold stale records of an earlier chunk count
        DeviceGuard guard_(hs[r]->device);
        for (int k = C; k < GSB_MAX_CHUNKS; ++k) CK(cudaEventRecord(hs[r]->ev_piece[k], hs[r]->s_comm));
        hs[r]->ghost_valid = true;
    }
    return 0;
}

// Array(fields.u) must show the neighbours' columns although, in the fused scheme, the kernels read
// them from the receive buffers: copy them into the ghost columns of the current fields on demand.
static int refresh_x_ghosts(gsb_handle* h) {
    if (!h->n_active || !h->ghost_valid || !fused_x(h)) return 0;
    const int C = pipeline_chunks(h);
    for (int k = 0; k < C; ++k) {
        Piece pc = chunk_piece(h, k, C);
        for (int d = 2; d < 6; ++d) pc.recv[d] = false;
        // unpack only (no neighbour wait: the data landed before ghost_valid was set; callers synchronised s_comm)
        FaceSet fs;
        fs.n = 0;
        for (int d = 0; d < 2; ++d) {
            if (!pc.recv[d]) continue;
            fs.dir[fs.n] = d;
            fs.buf[fs.n] = (char*)h->recv[h->step & 1][d] + piece_offset(h, d, pc.z0);
            fs.z0[fs.n] = pc.z0; fs.z1[fs.n] = pc.z1;
            fs.n++;
        }
        if (fs.n) CK(launch_unpack_faces(h->cfg.dtype, h->f[0], h->f[1], h->L, fs, h->s_compute));
    }
    return 0;
}

static int pipeline_step(gsb_handle** hs, int n) {
    if (pipeline_prime(hs, n)) return 1;
    const int C = pipeline_chunks(hs[0]);
    int order[GSB_MAX_CHUNKS];
    chunk_order(C, order);
    Params ps[64];
    if (n > 64) return fail("too many in-process ranks");
    for (int r = 0; r < n; ++r) { ps[r] = make_params(hs[r]); hs[r]->piece_cursor = 0; hs[r]->pieces_timed = 0; }
    for (int i = 0; i < C; ++i) {
        const int k = order[i];
        for (int r = 0; r < n; ++r) {
            gsb_handle* h = hs[r];
            if (pipeline_compute_chunk(h, k, C, ps[r], i)) return 1;
            if (h->n_active)
                if (enqueue_pack_send(h, chunk_piece(h, k, C), 1, (int)((h->step + 1) & 1), h->ev_chunk, fused_x(h))) return 1;
        }
        for (int r = 0; r < n; ++r) {
            gsb_handle* h = hs[r];
            if (h->n_active)
                if (enqueue_unpack(h, chunk_piece(h, k, C), 1, (int)((h->step + 1) & 1), h->ev_piece[k], fused_x(h))) return 1;
        }
    }
    for (int r = 0; r < n; ++r) { swap_bindings(hs[r]); hs[r]->ghost_valid = hs[r]->n_active > 0; }
    return 0;
}

static int split_step(gsb_handle** hs, int n) {
    for (int r = 0; r < n; ++r) { hs[r]->piece_cursor = 0; hs[r]->pieces_timed = 0; }
    for (int r = 0; r < n; ++r) if (hs[r]->n_active) if (enqueue_exchange_now(hs[r], false)) return 1;
    for (int r = 0; r < n; ++r) if (hs[r]->n_active) if (enqueue_exchange_now(hs[r], true)) return 1;
    for (int r = 0; r < n; ++r) if (enqueue_compute_overlapped(hs[r])) return 1;
    for (int r = 0; r < n; ++r) { swap_bindings(hs[r]); hs[r]->ghost_valid = false; }
    return 0;
}

extern "C" int gsb_exchange(gsb_t* h) {
    if (!h) return fail("null handle");
    if (!h->n_active) return 0;
    if (check_transport(h)) return 1;
    if (h->transport == TR_LOCAL) return fail("gsb_exchange: in-process handles advance together, use gsb_step_group");
    h->piece_cursor = 0; h->pieces_timed = 0;
    if (enqueue_exchange_now(h, false)) return 1;
    if (enqueue_exchange_now(h, true)) return 1;
    h->ghost_valid = false;   // whole-face message layout: a later gsb_step exchanges again in its own layout
    return 0;
}

extern "C" int gsb_calculate(gsb_t* h) {
    if (!h) return fail("null handle");
    DeviceGuard guard_(h->device);
    if (guard_.err != cudaSuccess) return fail("cudaSetDevice(%d) -> %s", h->device, cudaGetErrorString(guard_.err));
    if (ensure_maps(h)) return 1;
    const Params p = make_params(h);
    if (h->n_active) {
        // everything enqueued on the communication stream so far (gsb_exchange's unpack, or the pieces a
        // pipelined gsb_step sent ahead) must have landed in the ghosts
        CK(cudaEventRecord(h->ev_halo, h->s_comm));
        CK(cudaStreamWaitEvent(h->s_compute, h->ev_halo, 0));
        CK(cudaStreamWaitEvent(h->s_compute2, h->ev_halo, 0));
    }
    if (h->n_active && fused_x(h) && h->ghost_valid) if (refresh_x_ghosts(h)) return 1;
    if (launch_box(h, full_box(h), p)) return 1;
    CK(cudaEventRecord(h->ev_compute, h->s_compute));
    return 0;
}

extern "C" int gsb_swap(gsb_t* h) {
    if (!h) return fail("null handle");
    swap_bindings(h);
    h->ghost_valid = false;
    return 0;
}

extern "C" int gsb_step(gsb_t* h, int nsteps) {
    if (!h) return fail("null handle");
    if (nsteps < 0) return fail("nsteps < 0");
    if (check_transport(h)) return 1;
    if (h->transport == TR_LOCAL) return fail("gsb_step: in-process handles advance together, use gsb_step_group");
    for (int s = 0; s < nsteps; ++s)
        if (h->halo_mode == HALO_PIPELINE ? pipeline_step(&h, 1) : split_step(&h, 1)) return 1;
    return 0;
}

extern "C" int gsb_step_group(gsb_t** hs, int n, int nsteps) {
    if (!hs || n < 1) return fail("gsb_step_group: bad arguments");
    if (nsteps < 0) return fail("nsteps < 0");
    for (int r = 0; r < n; ++r) {
        if (!hs[r]) return fail("null handle %d", r);
        if (check_transport(hs[r])) return 1;
        if (hs[r]->n_active && hs[r]->transport != TR_LOCAL) return fail("gsb_step_group needs gsb_comm_init_local");
        if (hs[r]->step != hs[0]->step) return fail("handles are at different steps");
        if (hs[r]->L.sz != hs[0]->L.sz || hs[r]->halo_mode != hs[0]->halo_mode || hs[r]->halo_chunks != hs[0]->halo_chunks)
            return fail("in-process ranks must share sz and the halo scheme");
    }
    for (int s = 0; s < nsteps; ++s)
        if (hs[0]->halo_mode == HALO_PIPELINE ? pipeline_step(hs, n) : split_step(hs, n)) return 1;
    return 0;
}

// ------------------------------------------------------------------ stateless kernel entry
extern "C" int gsb_calculate_arrays(int dtype, int arith, const int64_t sizes[3], const int64_t offsets[3],
                                    const void* u_dev, const void* v_dev, void* u_out_dev, void* v_out_dev,
                                    double Du, double Dv, double F, double k, double dt, double noise,
                                    uint64_t seed, uint32_t step, void* stream) {
    if (!sizes || !offsets || !u_dev || !v_dev || !u_out_dev || !v_out_dev) return fail("gsb_calculate_arrays: null argument");
    if (dtype != GSB_F32 && dtype != GSB_F64) return fail("dtype must be GSB_F32 or GSB_F64");
    if (arith == GSB_ARITH_NATIVE && dtype != GSB_F32) return fail("GSB_ARITH_NATIVE exists for Float32 only");
    for (int d = 0; d < 3; ++d)
        if (sizes[d] < 1 || sizes[d] > (1 << 30)) return fail("sizes[%d] out of range", d);
    if (sizes[2] > 65535) return fail("sizes[2] > 65535 is not supported by this entry point");
    gsb_handle tmp;
    memset((void*)&tmp, 0, sizeof tmp);
    tmp.cfg.dtype = dtype; tmp.cfg.Du = Du; tmp.cfg.Dv = Dv; tmp.cfg.F = F; tmp.cfg.k = k; tmp.cfg.dt = dt;
    tmp.cfg.noise = noise; tmp.cfg.seed = seed; tmp.step = step;
    for (int d = 0; d < 3; ++d) tmp.cfg.offsets[d] = offsets[d];
    const Params p = make_params(&tmp);
    CK(launch_naive_reflayout(dtype, arith, noise != 0.0, u_dev, v_dev, u_out_dev, v_out_dev, (int)sizes[0],
                              (int)sizes[1], (int)sizes[2], p, (cudaStream_t)stream));
    return 0;
}
