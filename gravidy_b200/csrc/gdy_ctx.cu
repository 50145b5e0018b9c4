// gdy_ctx.cu — the context behind include/gravidy_b200.h: device memory, launch logic, NCCL plumbing
// and the C ABI itself.  sm_100a only; there is no CPU path in this file or anywhere in the library.
//
// One gdy_ctx = one GPU = one j-shard.  State is replicated on every shard; the force kernels of a
// shard see the j-tiles [tile_begin, tile_end) only, and the partial forces of the (shared) active list
// are summed over the shards with ncclAllReduce (reference's equivalent: the host loop
// h_f[id] += h_fout_gpu[g][i], src/include/gpu/Hermite4GPU.cu:512-530; MPI back end:
// src/include/mpi/Hermite4MPI.cpp:202).
#include "../../include/gravidy_b200.h"
#include "gdy_force.cuh"
#include "gdy_step.cuh"
#include "gdy_persist.cuh"
#include "gdy_persist2.cuh"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <future>
#include <memory>
#include <thread>
#include <unistd.h>
#include <nccl.h>   // types only; the library is bound with dlopen so that a process which already
                    // carries an NCCL (e.g. torch's bundled one) shares it instead of loading a second.
#include <vector>

using namespace gdy;

// ---- NCCL, bound at run time ---------------------------------------------------------------------------
namespace {
struct NcclApi
{
    void *h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};
NcclApi g_nccl;
char g_create_err[512] = "";

bool nccl_load()
{
    if (g_nccl.ok) return true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names)
    {
        g_nccl.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.h) break;
    }
    if (!g_nccl.h) return false;
#define GDY_SYM(field, name) \
    *(void **)(&g_nccl.field) = dlsym(g_nccl.h, name); \
    if (!g_nccl.field) return false;
    GDY_SYM(GetUniqueId, "ncclGetUniqueId")
    GDY_SYM(CommInitRank, "ncclCommInitRank")
    GDY_SYM(CommDestroy, "ncclCommDestroy")
    GDY_SYM(AllReduce, "ncclAllReduce")
    GDY_SYM(AllGather, "ncclAllGather")
    GDY_SYM(GetErrorString, "ncclGetErrorString")
#undef GDY_SYM
    g_nccl.ok = true;
    return true;
}
}  // namespace

// ---- the context ---------------------------------------------------------------------------------------
struct gdy_ctx
{
    int device = 0, rank = 0, world = 1;
    uint32_t n = 0, npad = 0;
    double e2 = 0.0;
    float eta = 0.0f;
    bool mask_self = false;          // e2 == 0: the self pair must be masked
    cudaStream_t st = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    ncclComm_t comm = nullptr;

    double *pool = nullptr;          // every double array of State lives here
    size_t pool_doubles = 0;
    State s{};
    uint32_t *d_move = nullptr;      // s.move
    uint32_t *d_move_alt = nullptr;  // the list gdy_next_integration_time prepares for the next find_particles_to_move
    uint32_t *d_ilist = nullptr;     // explicit list for gdy_force_sweep
    uint32_t *d_block_cnt = nullptr;
    uint32_t n_sched_blocks = 0;
    double *d_block_min = nullptr;
    uint32_t n_min_blocks = 0;
    SchedOut *d_sched = nullptr;
    double *d_atime = nullptr;
    double *d_energy = nullptr;      // [2]
    double *d_stage = nullptr;       // AoS staging, npad*8 doubles
    size_t part_cap = 0;             // doubles in s.part

    struct Pinned { SchedOut sched; double atime; double energy[2]; uint32_t abort; uint32_t atime_seq; } *pin = nullptr;

    uint32_t blk_begin = 0, blk_end = 0;     // this shard's j-blocks (units of JB = 64 particles)
    int num_sms = 148;
    int occ[FV_COUNT] = {};                  // resident CTAs per SM of each force variant
    uint32_t *d_counter = nullptr;           // arrival counters of the fused last-block reduction
    int tune_variant = -1, tune_jsplit = -1;
    uint32_t seq = 0;
    uint32_t last_nact = 0;
    // hook loop: next_integration_time already ran the scheduler for the time it returned (one wait per block step)
    bool sched_cached = false;
    double sched_cached_time = 0.0;
    bool have_particles = false, have_pred = false;
    uint64_t launches = 0;
    double force_ms_acc = 0.0;

    // persistent small-step kernel (gdy_persist2.cuh)
    uint32_t ps_nact_max = 0;        // 0 = off
    uint32_t ps_grid = 0, ps_gpad = 0;
    PersistCtl *ps_ctl = nullptr, *ps_ctl_host = nullptr;
    TraceRec *ps_trace_host = nullptr;
    uint32_t ps_seq = 0, ps_xseq = 0;
    bool ps_last = false;            // the last block step was done by the persistent kernel
    // its flagged-slot buffers and the shared-memory carve-up
    bool p2_ready = false;
    uint4 *p2_ipub = nullptr, *p2_part = nullptr, *p2_anext = nullptr, *p2_cmin = nullptr;
    uint2 *p2_ccnt = nullptr, *p2_clist = nullptr, *p2_ecnt = nullptr;
    uint32_t *p2_abort = nullptr;
    unsigned long long *p2_prof = nullptr;
    uint32_t p2_off_dyn = 0, p2_lf_pad = 0, p2_res_cap = 0, p2_ls_cap = 0, p2_smem = 0;
    unsigned long long spin_timeout_ns = 2000000000ull;   // watchdog of every in-kernel poll (GDY_SPIN_TIMEOUT_MS)
    // peer-memory exchange of the persistent path (world > 1)
    bool peers_ok = false;
    uint4 *xbuf = nullptr;           // my exchange buffer (peers store into it)
    uint4 *xslot[PS_MAX_WORLD] = {}; // every rank's buffer as seen from this device
    void *xopened[PS_MAX_WORLD] = {};// mappings opened with cudaIpcOpenMemHandle (other processes)
    uint64_t ps_steps_total = 0, ps_launches = 0;
    uint32_t xmid_seq = 0;           // exchanges of the mid-block peer path so far (tag and slot parity)
    bool sharded_upload = true;      // GDY_SHARDED_UPLOAD=0: every rank copies the whole particle set over PCIe
    bool use_xmid = true;            // GDY_XMID=0: blocks above the persistent cap go through ncclAllReduce whatever their size
    uint64_t xmid_count = 0, nccl_count = 0;
    char err[512] = "";
};

namespace {

int fail(gdy_ctx *c, int code, const char *fmt, ...)
{
    char *dst = c ? c->err : g_create_err;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
    return code;
}

#define CU(call) \
    do { \
        cudaError_t e_ = (call); \
        if (e_ != cudaSuccess) \
            return fail(c, GDY_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
    } while (0)

#define NC(call) \
    do { \
        ncclResult_t r_ = (call); \
        if (r_ != ncclSuccess) \
            return fail(c, GDY_ERR_NCCL, "%s:%d %s: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_)); \
    } while (0)

#define LAUNCHED(c) \
    do { \
        ++(c)->launches; \
        CU(cudaGetLastError()); \
    } while (0)

inline uint32_t cdiv(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
inline uint32_t round_up(uint32_t a, uint32_t b) { return cdiv(a, b) * b; }

int bind_device(gdy_ctx *c) { CU(cudaSetDevice(c->device)); return 0; }

// the watchdog of the peer-memory exchange (k_xrecv / k_correct) reports through pinned memory; every call that
// has just synchronised the stream looks at it
int check_abort(gdy_ctx *c)
{
    if (c->pin && c->pin->abort)
        return fail(c, GDY_ERR_STATE, "peer exchange watchdog: rank %d waited %.0f ms for a peer's partial forces "
                    "(a peer rank died, was never attached or runs a different schedule); the state is undefined",
                    c->rank, c->spin_timeout_ns * 1e-6);
    return 0;
}
#define BIND(c) do { int r_ = bind_device(c); if (r_) return r_; } while (0)

}  // namespace

// ---- lifetime ------------------------------------------------------------------------------------------
extern "C" int gdy_abi_version(void) { return 3; }

extern "C" int gdy_device_count(void)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess) return 0;
    int ok = 0;
    for (int d = 0; d < ndev; ++d)
    {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major >= 10) ++ok;
    }
    return ok;
}

extern "C" const char *gdy_last_error(const gdy_ctx *c) { return c ? c->err : g_create_err; }

extern "C" int gdy_nccl_unique_id(uint8_t *uid_out)
{
    gdy_ctx *c = nullptr;
    if (!uid_out) return fail(c, GDY_ERR_INVALID, "uid_out is NULL");
    if (!nccl_load()) return fail(c, GDY_ERR_NCCL, "libnccl.so.2 not loadable: %s", dlerror());
    static_assert(sizeof(ncclUniqueId) == GDY_NCCL_UID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    NC(g_nccl.GetUniqueId(&id));
    memcpy(uid_out, &id, sizeof id);
    return GDY_OK;
}

namespace {

constexpr size_t PART_MIN_DOUBLES = (size_t)8 << 20;   // 64 MiB of j-split partials at least
// exchange buffer of a rank (peers store into it): the persistent kernel's slots, then the mid-block region
constexpr size_t XBUF_SLOTS = (size_t)2 * PS_MAX_WORLD * PS_NACT_CAP * 6;
constexpr size_t XMID_SLOTS = (size_t)2 * XMID_MAXW * 6 * XMID_LD;
static_assert(XMID_MAXW == PS_MAX_WORLD, "one world limit");
constexpr uint32_t PS_STEPS_MAX = 4096;                // block steps per persistent launch
constexpr uint32_t PS_DEFAULT_NACT = 512;              // active sets up to this size go to the persistent kernel

// doubles of j-split partial sums: 64 MiB at least, and room for >= 4 j-splits of a full-active sweep
size_t part_cap_for(size_t np)
{
    return 4 * 6 * np > PART_MIN_DOUBLES ? 4 * 6 * np : PART_MIN_DOUBLES;
}

int alloc_all(gdy_ctx *c)
{
    const size_t np = c->npad;
    // 7 (r,m,v) + 6 f + 6 old + 3 a2 + 3 a3 + 2 (t,dt) + 7 pb + 6 fsum
    const size_t narr = 7 + 6 + 6 + 3 + 3 + 2 + 7 + 6;
    c->part_cap = part_cap_for(np);
    c->pool_doubles = narr * np + c->part_cap;
    CU(cudaMalloc(&c->pool, c->pool_doubles * sizeof(double)));
    CU(cudaMemsetAsync(c->pool, 0, c->pool_doubles * sizeof(double), c->st));
    double *q = c->pool;
    auto take = [&](size_t cnt) { double *r = q; q += cnt; return r; };
    State &s = c->s;
    s.x = take(np); s.y = take(np); s.z = take(np); s.m = take(np);
    s.vx = take(np); s.vy = take(np); s.vz = take(np);
    for (int k = 0; k < 6; ++k) s.f[k] = take(np);
    for (int k = 0; k < 6; ++k) s.old[k] = take(np);
    for (int k = 0; k < 3; ++k) s.a2[k] = take(np);
    for (int k = 0; k < 3; ++k) s.a3[k] = take(np);
    s.t = take(np); s.dt = take(np);
    s.pb = take(7 * np);          // 128-byte aligned: every array above is a multiple of 256 doubles
    s.fsum = take(6 * np);
    s.part = take(c->part_cap);

    CU(cudaMalloc(&c->d_move, np * sizeof(uint32_t)));
    CU(cudaMemsetAsync(c->d_move, 0xFF, np * sizeof(uint32_t), c->st));
    CU(cudaMalloc(&c->d_move_alt, np * sizeof(uint32_t)));
    CU(cudaMemsetAsync(c->d_move_alt, 0xFF, np * sizeof(uint32_t), c->st));
    s.move = c->d_move;
    CU(cudaMalloc(&c->d_ilist, np * sizeof(uint32_t)));
    CU(cudaMalloc(&c->d_counter, np * sizeof(uint32_t)));
    CU(cudaMemsetAsync(c->d_counter, 0, np * sizeof(uint32_t), c->st));
    s.tile_counter = c->d_counter;
    c->n_sched_blocks = cdiv(c->n, SCHED_CHUNK);
    CU(cudaMalloc(&c->d_block_cnt, c->n_sched_blocks * sizeof(uint32_t)));
    c->n_min_blocks = cdiv(c->n, 256 * 4);
    if (c->n_min_blocks > (uint32_t)MIN_BLOCKS) c->n_min_blocks = MIN_BLOCKS;
    CU(cudaMalloc(&c->d_block_min, MIN_BLOCKS * sizeof(double)));
    CU(cudaMalloc(&c->d_sched, sizeof(SchedOut)));
    CU(cudaMalloc(&c->d_atime, sizeof(double)));
    CU(cudaMalloc(&c->d_energy, 2 * sizeof(double)));
    CU(cudaMalloc(&c->d_stage, (np + 64) * 8 * sizeof(double)));   // + room for the padded chunks of the sharded upload
    CU(cudaHostAlloc(&c->pin, sizeof(*c->pin), cudaHostAllocDefault));
    memset(c->pin, 0, sizeof(*c->pin));

    // padding particles: zero mass, far away, at rest (they contribute exactly 0 to every sum)
    const uint32_t npd = c->npad - c->n;
    if (npd)
    {
        double *far[3] = {s.x, s.y, s.z};
        for (double *a : far)
        {
            k_fill<<<cdiv(npd, 256), 256, 0, c->st>>>(a, PAD_POS, c->n, c->npad);
            LAUNCHED(c);
        }
    }
    k_init_pb<<<cdiv(c->npad, 256), 256, 0, c->st>>>(s, c->n, c->npad);
    LAUNCHED(c);
    CU(force_variants_prepare());
    // persistent small-step kernel: one CTA per SM, co-resident (cooperative launch); its buffers and shared-memory
    // carve-up follow once the shard range is known (persist2_prepare)
    {
        int coop = 0;
        CU(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device));
        c->ps_grid = (uint32_t)c->num_sms;
        // GDY_PS_GRID: fewer CTAs than SMs, so that two logical ranks can run their persistent kernels side by side
        // on ONE device (tests of the peer-memory exchange on a single-GPU box)
        if (const char *e = getenv("GDY_PS_GRID")) { const int g = atoi(e); if (g >= 1 && g <= c->num_sms) c->ps_grid = (uint32_t)g; }
        if (c->ps_grid > (uint32_t)PS_GMAX) c->ps_grid = PS_GMAX;
        if (const char *e = getenv("GDY_XMID")) c->use_xmid = atoi(e) != 0;
        if (const char *e = getenv("GDY_SHARDED_UPLOAD")) c->sharded_upload = atoi(e) != 0;
        if (const char *e = getenv("GDY_SPIN_TIMEOUT_MS")) { const long ms = atol(e); if (ms > 0) c->spin_timeout_ns = (unsigned long long)ms * 1000000ull; }
        c->ps_gpad = round_up(c->ps_grid, 32);
        CU(cudaMalloc(&c->ps_ctl, sizeof(PersistCtl)));
        CU(cudaHostAlloc(&c->ps_ctl_host, sizeof(PersistCtl), cudaHostAllocDefault));
        CU(cudaHostAlloc(&c->ps_trace_host, PS_STEPS_MAX * sizeof(TraceRec), cudaHostAllocDefault));
        memset(c->ps_ctl_host, 0, sizeof(PersistCtl));
        uint32_t def = PS_DEFAULT_NACT;
        if (const char *e = getenv("GDY_PERSIST_NACT")) def = (uint32_t)atoi(e);
        if (def > (uint32_t)PS_NACT_CAP) def = PS_NACT_CAP;
        if (!coop) def = 0;
        // sharded contexts join the persistent path once their peers are attached (gdy_peer_attach)
        c->ps_nact_max = c->world == 1 ? def : 0;
    }
    for (int v = 0; v < FV_COUNT; ++v)
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->occ[v], (const void *)g_variants[v].fn[0],
                                                         g_variants[v].threads, g_variants[v].smem));
    CU(cudaStreamSynchronize(c->st));
    return 0;
}

}  // namespace

extern "C" int gdy_create_sharded(gdy_ctx **out, uint32_t n, double e2, float eta, int device, int rank,
                                  int world, const uint8_t *nccl_uid)
{
    gdy_ctx *c = nullptr;
    if (!out) return fail(c, GDY_ERR_INVALID, "out is NULL");
    *out = nullptr;
    if (n == 0) return fail(c, GDY_ERR_INVALID, "n must be > 0");
    // the warp force kernel's pinned instruction groups assume j indices below 2^26 (gdy_force.cuh, `opq`): above
    // that its never-taken branches WOULD be taken and the forces would be wrong without any error
    if (n > GDY_MAX_N) return fail(c, GDY_ERR_INVALID, "n = %u is above GDY_MAX_N = %u (the force kernel's index range)", n, (unsigned)GDY_MAX_N);
    if (!(e2 >= 0.0)) return fail(c, GDY_ERR_INVALID, "e2 must be >= 0");
    if (world < 1 || rank < 0 || rank >= world) return fail(c, GDY_ERR_INVALID, "bad rank/world %d/%d", rank, world);
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(c, GDY_ERR_NO_DEVICE, "no CUDA device (%s); gravidy_b200 has no CPU path",
                    ce == cudaSuccess ? "device count 0" : cudaGetErrorString(ce));
    if (device < 0 || device >= ndev) return fail(c, GDY_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);

    gdy_ctx *ctx = new gdy_ctx();
    ctx->device = device; ctx->rank = rank; ctx->world = world;
    ctx->n = n; ctx->npad = round_up(n, PAD);
    ctx->e2 = e2; ctx->eta = eta; ctx->mask_self = (e2 == 0.0);
    c = ctx;
    int rc = 0;
    do {
        if ((rc = bind_device(c))) break;
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { rc = fail(c, GDY_ERR_CUDA, "cudaGetDeviceProperties failed"); break; }
        if (prop.major < 10) { rc = fail(c, GDY_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor); break; }
        c->num_sms = prop.multiProcessorCount;
        auto cu = [&](cudaError_t e, const char *what) { if (e != cudaSuccess && !rc) rc = fail(c, GDY_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e)); };
        cu(cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking), "cudaStreamCreate");
        cu(cudaEventCreate(&c->ev0), "cudaEventCreate");
        cu(cudaEventCreate(&c->ev1), "cudaEventCreate");
        if (rc) break;
        if ((rc = alloc_all(c))) break;
        // contiguous, tile-aligned j-shards (reference partition: src/include/gpu/Hermite4GPU.cu:81-97)
        const uint32_t nblk = c->npad / JB;   // same arithmetic as gdy_shard_range
        c->blk_begin = (uint32_t)((uint64_t)nblk * rank / world);
        c->blk_end = (uint32_t)((uint64_t)nblk * (rank + 1) / world);
        if (world > 1 && nccl_uid)
        {
            if (!nccl_load()) { rc = fail(c, GDY_ERR_NCCL, "libnccl.so.2 not loadable: %s", dlerror()); break; }
            ncclUniqueId id;
            memcpy(&id, nccl_uid, sizeof id);
            // ncclCommInitRank blocks until all `world` ranks have called it: a rank that never comes would hang this
            // one for ever.  The call runs on a helper thread and this one waits GDY_NCCL_INIT_TIMEOUT_MS (default
            // 120 000) for it; on expiry the context is refused with an error (the reference's convention is report +
            // exit: src/include/gpu/Hermite4GPU.cuh:43-50) and the helper, which cannot be cancelled, is left behind.
            struct InitJob { ncclComm_t comm = nullptr; ncclResult_t res = ncclSuccess; std::promise<void> done; };
            std::shared_ptr<InitJob> job = std::make_shared<InitJob>();
            std::future<void> fut = job->done.get_future();
            std::thread([job, id, world, rank, device]() {
                cudaSetDevice(device);
                job->res = g_nccl.CommInitRank(&job->comm, world, id, rank);
                job->done.set_value();
            }).detach();
            long tmo = 120000;
            if (const char *e = getenv("GDY_NCCL_INIT_TIMEOUT_MS")) { const long v = atol(e); if (v > 0) tmo = v; }
            if (fut.wait_for(std::chrono::milliseconds(tmo)) != std::future_status::ready)
            {
                rc = fail(c, GDY_ERR_NCCL, "ncclCommInitRank: rank %d of %d waited %ld ms for the other ranks (a rank is missing or "
                          "holds a different unique id); the process must exit now: the init thread cannot be cancelled", rank, world, tmo);
                break;
            }
            if (job->res != ncclSuccess) { rc = fail(c, GDY_ERR_NCCL, "ncclCommInitRank: %s", g_nccl.GetErrorString(job->res)); break; }
            c->comm = job->comm;
        }
    } while (0);
    if (rc)
    {
        snprintf(g_create_err, sizeof g_create_err, "%s", c->err);
        gdy_destroy(c);
        return rc;
    }
    *out = c;
    return GDY_OK;
}

extern "C" int gdy_shard_range(uint32_t n, int rank, int world, uint32_t *begin_out, uint32_t *end_out)
{
    if (n == 0 || world < 1 || rank < 0 || rank >= world) return GDY_ERR_INVALID;
    const uint32_t npad = round_up(n, PAD), nblk = npad / JB;
    uint32_t b = (uint32_t)((uint64_t)nblk * rank / world) * JB;
    uint32_t e = (uint32_t)((uint64_t)nblk * (rank + 1) / world) * JB;
    if (b > n) b = n;
    if (e > n) e = n;
    if (begin_out) *begin_out = b;
    if (end_out) *end_out = e;
    return GDY_OK;
}

extern "C" int gdy_create(gdy_ctx **out, uint32_t n, double e2, float eta, int device)
{
    return gdy_create_sharded(out, n, e2, eta, device, 0, 1, nullptr);
}

extern "C" int gdy_destroy(gdy_ctx *c)
{
    if (!c) return GDY_OK;
    cudaSetDevice(c->device);
    if (c->st) cudaStreamSynchronize(c->st);
    if (c->comm && g_nccl.ok) g_nccl.CommDestroy(c->comm);
    cudaFree(c->pool); cudaFree(c->d_move); cudaFree(c->d_move_alt); cudaFree(c->d_ilist); cudaFree(c->d_block_cnt);
    cudaFree(c->d_block_min); cudaFree(c->d_sched); cudaFree(c->d_atime); cudaFree(c->d_energy);
    cudaFree(c->d_stage); cudaFree(c->d_counter);
    cudaFree(c->ps_ctl);
    cudaFree(c->p2_ipub); cudaFree(c->p2_part); cudaFree(c->p2_anext); cudaFree(c->p2_cmin); cudaFree(c->p2_ccnt);
    cudaFree(c->p2_clist); cudaFree(c->p2_ecnt); cudaFree(c->p2_abort); cudaFree(c->p2_prof);
    for (int r = 0; r < PS_MAX_WORLD; ++r)
        if (c->xopened[r]) cudaIpcCloseMemHandle(c->xopened[r]);
    cudaFree(c->xbuf);
    if (c->ps_ctl_host) cudaFreeHost(c->ps_ctl_host);
    if (c->ps_trace_host) cudaFreeHost(c->ps_trace_host);
    if (c->pin) cudaFreeHost(c->pin);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->st) cudaStreamDestroy(c->st);
    delete c;
    return GDY_OK;
}

// ---- state transfer ------------------------------------------------------------------------------------
namespace {

// host AoS (n x stride doubles) -> device columns.  cols[k] == NULL skips AoS column k.
int upload_cols(gdy_ctx *c, const double *host, int stride, int ncol, double *const *cols)
{
    Cols cc{};
    for (int k = 0; k < ncol; ++k) cc.c[k] = cols[k];
    CU(cudaMemcpyAsync(c->d_stage, host, (size_t)c->n * stride * sizeof(double), cudaMemcpyHostToDevice, c->st));
    k_unpack<<<cdiv(c->n, 256), 256, 0, c->st>>>(c->d_stage, cc, ncol, stride, c->n);
    LAUNCHED(c);
    CU(cudaStreamSynchronize(c->st));   // host buffer is borrowed for the call only
    return 0;
}

int download_cols(gdy_ctx *c, double *host, int stride, int ncol, double *const *cols)
{
    Cols cc{};
    for (int k = 0; k < ncol; ++k) cc.c[k] = cols[k];
    k_pack<<<cdiv(c->n, 256), 256, 0, c->st>>>(c->d_stage, cc, ncol, stride, c->n);
    LAUNCHED(c);
    CU(cudaMemcpyAsync(host, c->d_stage, (size_t)c->n * stride * sizeof(double), cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    return check_abort(c);
}

}  // namespace

extern "C" int gdy_upload_particles(gdy_ctx *c, const double *r4, const double *v4)
{
    if (!c) return GDY_ERR_INVALID;
    if (!r4 || !v4) return fail(c, GDY_ERR_INVALID, "r4/v4 is NULL");
    BIND(c);
    c->sched_cached = false;
    State &s = c->s;
    Cols rc{}, vc{};
    rc.c[0] = s.x; rc.c[1] = s.y; rc.c[2] = s.z; rc.c[3] = s.m;
    vc.c[0] = s.vx; vc.c[1] = s.vy; vc.c[2] = s.vz;
    const size_t n = c->n;
    if (c->world > 1 && c->comm && c->sharded_upload)
    {
        // A sharded run uploads collectively: every rank is handed the same full arrays (the state is replicated), so a
        // rank copies only ITS chunk over PCIe and the chunks are all-gathered over NVLink (the reference pushes the
        // whole set to every GPU through the host link: src/include/gpu/Hermite4GPU.cu:226-246).
        const size_t chunk = cdiv(c->n, (uint32_t)c->world);
        const size_t b = chunk * c->rank < n ? chunk * c->rank : n, e = b + chunk < n ? b + chunk : n;
        double *mine = c->d_stage + (size_t)c->rank * chunk * 8;
        if (e > b)
        {
            CU(cudaMemcpyAsync(mine, r4 + b * 4, (e - b) * 4 * sizeof(double), cudaMemcpyHostToDevice, c->st));
            CU(cudaMemcpyAsync(mine + chunk * 4, v4 + b * 4, (e - b) * 4 * sizeof(double), cudaMemcpyHostToDevice, c->st));
        }
        NC(g_nccl.AllGather(mine, c->d_stage, chunk * 8, ncclDouble, c->comm, c->st));
        k_unpack_chunks<<<cdiv(c->n, 256), 256, 0, c->st>>>(c->d_stage, rc, vc, (uint32_t)chunk, c->n);
        LAUNCHED(c);
    }
    else
    {
        CU(cudaMemcpyAsync(c->d_stage, r4, n * 4 * sizeof(double), cudaMemcpyHostToDevice, c->st));
        CU(cudaMemcpyAsync(c->d_stage + n * 4, v4, n * 4 * sizeof(double), cudaMemcpyHostToDevice, c->st));
        k_unpack<<<cdiv(c->n, 256), 256, 0, c->st>>>(c->d_stage, rc, 4, 4, c->n);
        LAUNCHED(c);
        k_unpack<<<cdiv(c->n, 256), 256, 0, c->st>>>(c->d_stage + n * 4, vc, 4, 4, c->n);
        LAUNCHED(c);
    }
    k_init_pb<<<cdiv(c->npad, 256), 256, 0, c->st>>>(s, c->n, c->npad);   // mass row of the predicted set
    LAUNCHED(c);
    CU(cudaStreamSynchronize(c->st));   // the host buffers are borrowed for the call only: ONE synchronisation
    c->have_particles = true;
    c->have_pred = false;
    return GDY_OK;
}

extern "C" int gdy_upload_times(gdy_ctx *c, const double *t, const double *dt)
{
    if (!c) return GDY_ERR_INVALID;
    if (!t || !dt) return fail(c, GDY_ERR_INVALID, "t/dt is NULL");
    BIND(c);
    c->sched_cached = false;
    CU(cudaMemcpyAsync(c->s.t, t, (size_t)c->n * sizeof(double), cudaMemcpyHostToDevice, c->st));
    CU(cudaMemcpyAsync(c->s.dt, dt, (size_t)c->n * sizeof(double), cudaMemcpyHostToDevice, c->st));
    CU(cudaStreamSynchronize(c->st));
    return GDY_OK;
}

extern "C" int gdy_upload_forces(gdy_ctx *c, const double *f6)
{
    if (!c) return GDY_ERR_INVALID;
    if (!f6) return fail(c, GDY_ERR_INVALID, "f6 is NULL");
    BIND(c);
    return upload_cols(c, f6, 6, 6, c->s.f);
}

extern "C" int gdy_upload_derivs(gdy_ctx *c, const double *old6, const double *a2, const double *a3)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    int r;
    if (old6 && (r = upload_cols(c, old6, 6, 6, c->s.old))) return r;
    if (a2 && (r = upload_cols(c, a2, 3, 3, c->s.a2))) return r;
    if (a3 && (r = upload_cols(c, a3, 3, 3, c->s.a3))) return r;
    return GDY_OK;
}

extern "C" int gdy_download_forces(gdy_ctx *c, double *f6)
{
    if (!c) return GDY_ERR_INVALID;
    if (!f6) return fail(c, GDY_ERR_INVALID, "f6 is NULL");
    BIND(c);
    return download_cols(c, f6, 6, 6, c->s.f);
}

extern "C" int gdy_download_state(gdy_ctx *c, double *r4, double *v4, double *f6, double *old6, double *a2,
                                  double *a3, double *t, double *dt, double *p7, uint32_t *move)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    State &s = c->s;
    int r;
    double *rc[4] = {s.x, s.y, s.z, s.m};
    double *vc[4] = {s.vx, s.vy, s.vz, nullptr};
    if (r4 && (r = download_cols(c, r4, 4, 4, rc))) return r;
    if (v4 && (r = download_cols(c, v4, 4, 4, vc))) return r;
    if (f6 && (r = download_cols(c, f6, 6, 6, s.f))) return r;
    if (old6 && (r = download_cols(c, old6, 6, 6, s.old))) return r;
    if (a2 && (r = download_cols(c, a2, 3, 3, s.a2))) return r;
    if (a3 && (r = download_cols(c, a3, 3, 3, s.a3))) return r;
    if (p7)
    {
        k_pack_pred<<<cdiv(c->n, 256), 256, 0, c->st>>>(c->d_stage, s, c->n);
        LAUNCHED(c);
        CU(cudaMemcpyAsync(p7, c->d_stage, (size_t)c->n * 7 * sizeof(double), cudaMemcpyDeviceToHost, c->st));
        CU(cudaStreamSynchronize(c->st));
    }
    if (t) CU(cudaMemcpyAsync(t, s.t, (size_t)c->n * sizeof(double), cudaMemcpyDeviceToHost, c->st));
    if (dt) CU(cudaMemcpyAsync(dt, s.dt, (size_t)c->n * sizeof(double), cudaMemcpyDeviceToHost, c->st));
    if (move) CU(cudaMemcpyAsync(move, s.move, (size_t)c->n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    if (move)   // entries past the active count read -1, as the reference leaves them (Hermite4.cpp:69)
        for (uint32_t k = c->last_nact; k < c->n; ++k) move[k] = 0xFFFFFFFFu;
    return GDY_OK;
}

// ---- the force path --------------------------------------------------------------------------------------
namespace {

struct ForcePlan
{
    int variant;
    dim3 grid;
    uint32_t jb, je, per_split, nsplit, ld;
};

// Which kernel for how many active particles (set from tools/sweep_force.py on a B200, see
// profiles/): small blocks put j across lanes and read it straight from L2; from a few hundred active
// particles on, i sits in registers and j streams through the per-warp TMA rings.
int pick_variant(const gdy_ctx *c, uint32_t nact)
{
    if (c->tune_variant >= 0 && c->tune_variant < FV_COUNT) return c->tune_variant;
    if (nact <= 2) return FV_LANE_I1x2;
    const uint64_t nj = (uint64_t)(c->blk_end - c->blk_begin) * JB;
    // j across lanes straight from L2 has the lower fixed cost (no pipeline fill, no second launch);
    // it wins while the whole evaluation is a few tens of microseconds
    if (nact <= 160 || (uint64_t)nact * nj <= 12000000ull) return FV_LANE_I1x4;
    // A block that cannot fill two waves of 64-i tiles (few i-tiles x few j-blocks: the mid blocks of a sharded run,
    // 1172 active against an eighth of N = 256k) takes 32-i tiles: twice the CTAs to balance the SMs with
    // (profiles/r02_sweep_shard_w8.md: 94 instead of 114 us at nact 1172, 66 instead of 78 at 600, even at 2400, 1 % slower
    // from 4800 on)
    if ((uint64_t)nact * (c->blk_end - c->blk_begin) < 1000000ull) return FV_WARP_I32;
    // the smallest i-tile of the two-i-per-lane family at every other size: with ~24 waves of small CTAs (plan_force)
    // the block scheduler balances the SMs dynamically, and a CTA's prologue hides behind its neighbours' FP64 work
    // (profiles/archive/r01_sweep_split_v11.md: 517 vs 508 Gpairs/s for warp_i256 at full-active N = 256k)
    return FV_WARP_I64;
}

// j-split: CTAs = i-tiles x splits.  Lane kernels (latency-bound small blocks): the cost of a plan is
// waves x (j-units per CTA + a fixed per-CTA overhead) with waves of (SMs x resident CTAs); minimising it
// fills the machine.  Warp kernels: ~24 waves of small CTAs (see below).
ForcePlan plan_force(const gdy_ctx *c, uint32_t nact)
{
    ForcePlan p{};
    p.variant = pick_variant(c, nact);
    const ForceVariant &v = g_variants[p.variant];
    const uint32_t gx = cdiv(nact, (uint32_t)v.ti);
    const uint32_t slots = (uint32_t)c->num_sms * (uint32_t)(c->occ[p.variant] > 0 ? c->occ[p.variant] : 1);
    p.ld = round_up(nact, 32);
    uint32_t units;   // j work items that can be split: blocks of 64 (warp kernels) or trips (lane kernels)
    if (v.kind == 0)
    {
        p.jb = c->blk_begin; p.je = c->blk_end;
        units = p.je - p.jb;
    }
    else
    {
        p.jb = c->blk_begin * JB;
        p.je = c->blk_end * JB < c->n ? c->blk_end * JB : c->n;
        if (p.je < p.jb) p.je = p.jb;
        units = cdiv(p.je - p.jb, (uint32_t)v.pj);
    }
    uint32_t split = 1;
    if (c->tune_jsplit > 0) split = (uint32_t)c->tune_jsplit;
    else if (units > 1 && v.kind == 0)
    {
        // Warp kernels: many small CTAs.  Measured (tools/sweep_split.py, profiles/archive/r01_sweep_split_v11.md): the rate
        // keeps rising until ~20-28 waves of CTAs and down to ~16 j-blocks per CTA — the one-wave plans of the old
        // wave model lost 2-8 % at awkward sizes (an SM that is a little slower, or has 3 CTAs instead of 4,
        // decides the time), while per-CTA fixed costs overlap with the other resident CTAs' FP64 work.
        split = cdiv(24u * slots, gx);
        const uint32_t cap = units / 16 > 1 ? units / 16 : 1;
        if (split > 192) split = 192;
        if (split > cap) split = cap;
        // The last wave: CTAs of one launch take the same time, so the SMs run in step and what is left over for the
        // last round decides — 19 i-tiles x 187 splits are 3553 CTAs, 24 rounds of 148 and ONE CTA more (646 us where
        // 590 would do); 19 x 32 on an eighth of the j-particles are 608 CTAs on 592 slots.  The SM with the most CTAs
        // has ceil(CTAs / SMs) of them and the FP64 pipe makes their times add up, so of the splits between three
        // quarters of the target and the target take the one with the smallest  ceil(CTAs / SMs) x j-blocks per CTA  (a CTA alone on
        // its SM is latency-bound: x 1.7), the larger split on a tie (profiles/r02_sweep_shard_w8.md).
        {
            const uint32_t sms = (uint32_t)c->num_sms;
            uint32_t best = split;
            double best_cost = 1e300;
            for (uint32_t sp = split; sp >= 1 && 4 * sp >= 3 * split; --sp)
            {
                const uint32_t per = cdiv(units, sp), s2 = cdiv(units, per);
                if (s2 != sp) continue;   // the same plan as a smaller sp
                const uint32_t q = cdiv(gx * s2, sms);
                const double cost = (q == 1 ? 1.7 : (double)q) * ((double)per + 0.5);
                if (cost < best_cost - 1e-9) { best_cost = cost; best = sp; }
            }
            split = best;
        }
    }
    else if (units > 1)
    {
        const uint64_t overhead = v.kind == 0 ? 3 : 2;   // per-CTA fixed cost in j-units
        uint32_t smax = (8 * slots) / gx + 1;
        if (smax > units) smax = units;
        if (smax > 2048) smax = 2048;
        uint64_t best = ~0ull;
        for (uint32_t sp = 1; sp <= smax; ++sp)
        {
            const uint64_t waves = cdiv(gx * sp, slots);
            const uint64_t cost = waves * (cdiv(units, sp) + overhead);
            if (cost < best) { best = cost; split = sp; }
        }
    }
    const size_t max_split = c->part_cap / ((size_t)6 * p.ld);
    if (split > max_split) split = (uint32_t)max_split;
    if (split > units) split = units;
    if (split < 1) split = 1;
    uint32_t per = units ? cdiv(units, split) : 1;
    split = units ? cdiv(units, per) : 1;
    p.nsplit = split;
    p.per_split = v.kind == 0 ? per : per * (uint32_t)v.pj;
    p.grid = dim3(gx, split, 1);
    return p;
}

}  // namespace

// The launch plan of a force evaluation without touching a device (host logic only; tests/test_capi_cpu.py):
// out = {variant, kind (0 warp / 1 lane), grid.x = i-tiles, grid.y = j-splits, j-units per split, first j-unit,
// end j-unit, ld}; j-units are blocks of 64 particles for the warp kernels and particles for the lane kernels.
extern "C" int gdy_plan_dry(uint32_t n, uint32_t nact, int rank, int world, int num_sms, int resident_ctas,
                            uint32_t *out /* [8] */)
{
    if (!out || n == 0 || nact == 0 || nact > n || world < 1 || rank < 0 || rank >= world || num_sms < 1)
        return GDY_ERR_INVALID;
    gdy_ctx c;
    c.n = n; c.npad = round_up(n, PAD); c.rank = rank; c.world = world;
    const uint32_t nblk = c.npad / JB;
    c.blk_begin = (uint32_t)((uint64_t)nblk * rank / world);
    c.blk_end = (uint32_t)((uint64_t)nblk * (rank + 1) / world);
    c.num_sms = num_sms;
    for (int v = 0; v < FV_COUNT; ++v) c.occ[v] = resident_ctas > 0 ? resident_ctas : 4;
    c.part_cap = part_cap_for(c.npad);
    const ForcePlan p = plan_force(&c, nact);
    out[0] = (uint32_t)p.variant; out[1] = (uint32_t)g_variants[p.variant].kind;
    out[2] = p.grid.x; out[3] = p.grid.y; out[4] = p.per_split; out[5] = p.jb; out[6] = p.je; out[7] = p.ld;
    return GDY_OK;
}

namespace {

// Partial forces of `nact` i-particles (move == NULL: particles 0..nact-1) against this shard's
// j-particles, summed over the j-splits and over the shards.  *out = k-indexed [6][ld] result buffer.
// xr_out != NULL: where the ranks' sums travel through peer memory, the receive is left to the caller's corrector
// (*xr_out filled in, base != NULL); otherwise *out holds the world's sums when this returns.
int force_eval(gdy_ctx *c, uint32_t nact, const uint32_t *move, const double **out, uint32_t *ld_out, bool timed,
               XRecv *xr_out = nullptr)
{
    if (xr_out) *xr_out = XRecv{};
    const State &s = c->s;
    const ForcePlan p = plan_force(c, nact);
    const ForceVariant &v = g_variants[p.variant];
    ForceArgs a{};
    a.pb = s.pb; a.move = move; a.part = s.part; a.fsum = s.fsum; a.counter = s.tile_counter;
    a.nact = nact; a.ld = p.ld; a.jb = p.jb; a.je = p.je; a.per_split = p.per_split; a.e2 = c->e2;
    if (timed) CU(cudaEventRecord(c->ev0, c->st));
    v.fn[c->mask_self ? 1 : 0]<<<p.grid, v.threads, v.smem, c->st>>>(a);
    ++c->launches;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(c, GDY_ERR_CUDA, "force kernel %s grid (%u,%u): %s", v.name, p.grid.x, p.grid.y, cudaGetErrorString(e));
    const bool xmid = c->world > 1 && c->peers_ok && c->use_xmid && nact <= (uint32_t)XMID_LD;
    XSend xs{};
    if (xmid)
    {
        for (int r = 0; r < c->world; ++r) xs.peer[r] = c->xslot[r] + XBUF_SLOTS;
        xs.rank = c->rank; xs.world = c->world; xs.tag = ++c->xmid_seq; xs.par = xs.tag & 1u; xs.xld = XMID_LD;
    }
    double *buf;
    bool sent = false;
    if (v.kind == 1) buf = s.fsum;            // lane kernels reduce their own j-splits
    else if (p.nsplit > 1)
    {
        // the reduction over the j-splits and the send half of the exchange in one launch
        k_reduce_partials<<<dim3(cdiv(nact, 32), 6), 256, 0, c->st>>>(s.part, s.fsum, nact, p.ld, p.nsplit, xs);
        LAUNCHED(c);
        buf = s.fsum;
        sent = xmid;
    }
    else buf = s.part;
    if (timed) CU(cudaEventRecord(c->ev1, c->st));
    if (xmid)
    {
        if (!sent)
        {
            k_xsend<<<cdiv(nact, 256), 256, 0, c->st>>>(buf, p.ld, nact, xs);
            LAUNCHED(c);
        }
        XRecv xr{c->xbuf + XBUF_SLOTS, c->world, xs.tag, xs.par, &c->pin->abort, c->spin_timeout_ns};
        ++c->xmid_count;
        if (xr_out) *xr_out = xr;
        else
        {
            k_xrecv<<<cdiv(nact, 256), 256, 0, c->st>>>(s.fsum, p.ld, nact, xr);
            LAUNCHED(c);
            buf = s.fsum;
        }
    }
    else if (c->world > 1 && c->comm)
    {
        NC(g_nccl.AllReduce(buf, buf, (size_t)6 * p.ld, ncclDouble, ncclSum, c->comm, c->st));
        ++c->nccl_count;
    }
    else if (c->world > 1 && c->peers_ok)
        return fail(c, GDY_ERR_STATE, "a block of %u active particles needs the NCCL communicator (peer exchange holds %d)", nact, (int)XMID_LD);
    *out = buf;
    *ld_out = p.ld;
    return 0;
}

int need_particles(gdy_ctx *c)
{
    if (!c->have_particles) return fail(c, GDY_ERR_STATE, "no particles uploaded (gdy_upload_particles)");
    return 0;
}

}  // namespace

extern "C" int gdy_set_tuning(gdy_ctx *c, int variant, int jsplit)
{
    if (!c) return GDY_ERR_INVALID;
    if (variant >= FV_COUNT) return fail(c, GDY_ERR_INVALID, "variant %d out of range (< %d)", variant, (int)FV_COUNT);
    c->tune_variant = variant;
    c->tune_jsplit = jsplit;
    return GDY_OK;
}

extern "C" int gdy_init_acc_jrk(gdy_ctx *c)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    int r;
    if ((r = need_particles(c))) return r;
    State &s = c->s;
    k_copy_to_pred<<<cdiv(c->n, 256), 256, 0, c->st>>>(s, c->n);
    LAUNCHED(c);
    c->have_pred = true;
    const double *buf; uint32_t ld;
    if ((r = force_eval(c, c->n, nullptr, &buf, &ld, false))) return r;
    k_scatter_forces<<<cdiv(c->n, 256), 256, 0, c->st>>>(s, nullptr, c->n, buf, ld);
    LAUNCHED(c);
    return GDY_OK;
}

extern "C" int gdy_init_dt(gdy_ctx *c, float eta_s, double itime, double *atime_out)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    c->sched_cached = false;
    // Host arithmetic with libm log/pow/ceil exactly as the reference does it (src/include/Hermite4.cpp:
    // 108-140): a device log could round differently and flip a timestep level.
    std::vector<double> f((size_t)c->n * 6), t(c->n), dt(c->n);
    int r;
    if ((r = gdy_download_forces(c, f.data()))) return r;
    double ctime = 1.0e+10;   // reference: Hermite4CPU_integration.cpp:43
    for (uint32_t i = 0; i < c->n; ++i)
    {
        const double *fi = &f[(size_t)i * 6];
        const double amag = std::sqrt(fi[0] * fi[0] + fi[1] * fi[1] + fi[2] * fi[2]);
        const double jmag = std::sqrt(fi[3] * fi[3] + fi[4] * fi[4] + fi[5] * fi[5]);
        double step = eta_s * (amag / jmag);
        const int level = (int)(std::ceil(std::log(step) / std::log(2.0)) - 1);
        step = std::pow(2, level);
        if (step < D_TIME_MIN) step = D_TIME_MIN;
        else if (step > D_TIME_MAX) step = D_TIME_MAX;
        dt[i] = step;
        t[i] = itime;
        if (step < ctime) ctime = step;
    }
    if ((r = gdy_upload_times(c, t.data(), dt.data()))) return r;
    if (atime_out) *atime_out = ctime;
    return GDY_OK;
}

// ---- the per-step hooks ------------------------------------------------------------------------------------
namespace {

// find_particles_to_move on the device.  from_min: ITIME = the minimum left in d_block_min by
// k_min_time (chained scheduler), else the `itime` argument.  Leaves {itime, nact} in the pinned slot.
int launch_sched(gdy_ctx *c, double itime, bool from_min, uint32_t *list_out = nullptr)
{
    State s = c->s;
    if (list_out) s.move = list_out;   // the hook loop's look-ahead: State.move keeps the current step's list
    const double *bm = from_min ? c->d_block_min : nullptr;
    ++c->seq;
    k_sched_count<<<c->n_sched_blocks, SCHED_BLOCK, 0, c->st>>>(s, c->n, itime, bm, c->n_min_blocks, c->d_block_cnt);
    LAUNCHED(c);
    k_sched_write<<<c->n_sched_blocks, SCHED_BLOCK, 0, c->st>>>(s, c->n, itime, bm, c->n_min_blocks, c->d_block_cnt,
                                                               c->n_sched_blocks, c->d_sched, &c->pin->sched, c->seq);
    LAUNCHED(c);
    return 0;
}

// Wait until a kernel has written `want` into a pinned word (its last store, behind a __threadfence_system): the host
// polls the word instead of waiting for the stream to drain — what is still running behind the store (the other blocks
// writing the active list, the other CTAs of the persistent kernel leaving) is ordered before the next launches by the
// stream.  8 us per block step cheaper than cudaStreamSynchronize (r02s).  cudaStreamQuery every few hundred polls
// catches a failed launch (the word would never arrive).
int wait_pinned(gdy_ctx *c, const volatile uint32_t *word, uint32_t want, const char *what)
{
    for (uint32_t spin = 0; *word != want; ++spin)
    {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();   // a polite spin: the sibling hyper-thread (another rank's host thread) keeps its share
#endif
        if ((spin & 255u) == 255u)
        {
            const cudaError_t q = cudaStreamQuery(c->st);
            if (q == cudaErrorNotReady) continue;
            if (q != cudaSuccess) return fail(c, GDY_ERR_CUDA, "%s: %s", what, cudaGetErrorString(q));
            if (*word != want) return fail(c, GDY_ERR_STATE, "%s left no result (seq %u != %u)", what, *word, want);
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return 0;
}

// the scheduler's {ITIME, nact} of launch_sched (k_sched_write, block 0, as soon as the counts are summed)
int wait_sched(gdy_ctx *c) { return wait_pinned(c, &c->pin->sched.seq, c->seq, "the scheduler kernel"); }

int launch_min_time(gdy_ctx *c)
{
    k_min_time<<<c->n_min_blocks, 256, 0, c->st>>>(c->s, c->n, c->d_block_min);
    LAUNCHED(c);
    return 0;
}

// predicted_pos_vel for this shard's j-slice (world == 1: everybody) plus `nextra` active particles
int launch_predict(gdy_ctx *c, double itime, uint32_t nextra)
{
    uint32_t b = c->blk_begin * JB, e = c->blk_end * JB;
    if (e > c->n) e = c->n;
    if (e < b) e = b;
    if (c->world == 1) nextra = 0;   // the slice is everybody
    const uint32_t cnt = (e - b) + nextra;
    if (cnt)
    {
        k_predict<<<cdiv(cnt, 256), 256, 0, c->st>>>(c->s, b, e, itime, c->s.move, nextra);
        LAUNCHED(c);
    }
    c->have_pred = true;
    return 0;
}

}  // namespace

extern "C" int gdy_find_particles_to_move(gdy_ctx *c, double itime, uint32_t *nact_out)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    int r;
    if (c->sched_cached && itime == c->sched_cached_time)
    {
        // gdy_next_integration_time left the list (in the alternate buffer) and the count for exactly this time
        c->sched_cached = false;
        std::swap(c->d_move, c->d_move_alt);
        c->s.move = c->d_move;
    }
    else
    {
        c->sched_cached = false;
        if ((r = launch_sched(c, itime, false))) return r;
        if ((r = wait_sched(c))) return r;
        if ((r = check_abort(c))) return r;
    }
    c->last_nact = c->pin->sched.nact;
    if (nact_out) *nact_out = c->last_nact;
    return GDY_OK;
}

extern "C" int gdy_save_old_acc_jrk(gdy_ctx *c, uint32_t nact)
{
    if (!c) return GDY_ERR_INVALID;
    if (nact > c->n) return fail(c, GDY_ERR_INVALID, "nact %u > n %u", nact, c->n);
    BIND(c);
    if (nact == 0) return GDY_OK;
    k_save_old<<<cdiv(nact, 128), 128, 0, c->st>>>(c->s, nact);
    LAUNCHED(c);
    return GDY_OK;
}

extern "C" int gdy_predicted_pos_vel(gdy_ctx *c, double itime)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    int r;
    if ((r = need_particles(c))) return r;
    // hook semantics = all N particles predicted on every shard (the fused driver predicts only the
    // shard's slice plus the active particles)
    k_predict<<<cdiv(c->n, 256), 256, 0, c->st>>>(c->s, 0, c->n, itime, nullptr, 0);
    LAUNCHED(c);
    c->have_pred = true;
    return GDY_OK;
}

// The same launch bracketed by events: the predictor is the memory-bound kernel of the step (120 B read +
// 48 B written per particle), bench.py reports its GB/s against the measured HBM copy bandwidth.
extern "C" int gdy_predict_timed(gdy_ctx *c, double itime, float *ms_out)
{
    if (!c || !ms_out) return GDY_ERR_INVALID;
    BIND(c);
    int r;
    if ((r = need_particles(c))) return r;
    CU(cudaEventRecord(c->ev0, c->st));
    k_predict<<<cdiv(c->n, 256), 256, 0, c->st>>>(c->s, 0, c->n, itime, nullptr, 0);
    LAUNCHED(c);
    CU(cudaEventRecord(c->ev1, c->st));
    CU(cudaStreamSynchronize(c->st));
    CU(cudaEventElapsedTime(ms_out, c->ev0, c->ev1));
    c->have_pred = true;
    return GDY_OK;
}

extern "C" int gdy_update_acc_jrk(gdy_ctx *c, uint32_t nact)
{
    if (!c) return GDY_ERR_INVALID;
    if (nact > c->n) return fail(c, GDY_ERR_INVALID, "nact %u > n %u", nact, c->n);
    BIND(c);
    if (!c->have_pred) return fail(c, GDY_ERR_STATE, "update_acc_jrk before predicted_pos_vel/init_acc_jrk");
    if (nact == 0) return GDY_OK;
    State &s = c->s;
    const double *buf; uint32_t ld;
    int r;
    if ((r = force_eval(c, nact, s.move, &buf, &ld, false))) return r;
    k_scatter_forces<<<cdiv(nact, 128), 128, 0, c->st>>>(s, s.move, nact, buf, ld);
    LAUNCHED(c);
    return GDY_OK;
}

extern "C" int gdy_correction_pos_vel(gdy_ctx *c, double itime, uint32_t nact)
{
    if (!c) return GDY_ERR_INVALID;
    if (nact > c->n) return fail(c, GDY_ERR_INVALID, "nact %u > n %u", nact, c->n);
    BIND(c);
    c->sched_cached = false;
    if (nact == 0) return GDY_OK;
    k_correct<<<cdiv(nact, 128), 128, 0, c->st>>>(c->s, nact, itime, (double)c->eta, nullptr, 0, XRecv{});
    LAUNCHED(c);
    return GDY_OK;
}

extern "C" int gdy_next_integration_time(gdy_ctx *c, double *atime_out)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    int r;
    // The scheduler of the NEXT block step is chained behind the minimum (as in gdy_run): the loop's next call is
    // find_particles_to_move(ATIME) with the time returned here, which then has its answer already — one host wait per
    // block step of the hook loop instead of two.  Any call that changes t or dt in between drops the cached answer.
    if ((r = launch_min_time(c))) return r;
    if ((r = launch_sched(c, 0.0, true, c->d_move_alt))) return r;
    if ((r = wait_sched(c))) return r;
    if ((r = check_abort(c))) return r;
    c->sched_cached = true;
    c->sched_cached_time = c->pin->sched.itime;
    if (atime_out) *atime_out = c->pin->sched.itime;
    return GDY_OK;
}

// get_energy_gpu(): the i-chunks (512 particles) are dealt to the shards in contiguous runs, every shard sums its chunks
// over ALL j, the chunk sums travel in one allreduce whose other entries are zeros (x + 0 is exact, so every rank ends
// up with every chunk sum bit for bit), and the total is a fixed-order tree over the chunks: the same bits on 1, 2, 4
// or 8 GPUs (reference: i-slices per GPU, src/include/gpu/Hermite4GPU.cu:540-592).  A logical shard (world > 1 without a
// communicator) returns the sum over its own chunks only.
namespace {
int energy_launch(gdy_ctx *c, float *ms_out)
{
    State &s = c->s;
    const uint32_t nch_total = cdiv(c->n, EN_CHUNK), ntiles = c->npad / EN_TILE;
    const uint32_t ch0 = (uint32_t)((uint64_t)nch_total * c->rank / c->world);
    const uint32_t ch1 = (uint32_t)((uint64_t)nch_total * (c->rank + 1) / c->world);
    // j-splits: a function of n alone (the decomposition must not depend on the number of GPUs): ~8 CTAs per SM-slot
    uint32_t nsplit = cdiv(8u * 148u, nch_total);
    if (nsplit > ntiles / 4) nsplit = ntiles / 4 ? ntiles / 4 : 1;
    if (nsplit < 1) nsplit = 1;
    const uint32_t per = cdiv(ntiles, nsplit);
    nsplit = cdiv(ntiles, per);
    double *cta_ek = s.part, *cta_ep = s.part + nch_total, *sums = cta_ep + (size_t)nsplit * nch_total;   // scratch
    if ((size_t)(nsplit + 3) * nch_total > c->part_cap) return fail(c, GDY_ERR_STATE, "energy scratch too small");
    if (ms_out) CU(cudaEventRecord(c->ev0, c->st));
    if (ch1 > ch0)
    {
        k_energy<<<dim3(ch1 - ch0, nsplit), EN_T, 0, c->st>>>(s, c->n, ch0, per, ntiles, cta_ek, cta_ep, nch_total);
        LAUNCHED(c);
    }
    if (ms_out) CU(cudaEventRecord(c->ev1, c->st));
    k_energy_fold<<<cdiv(nch_total, 256), 256, 0, c->st>>>(cta_ek, cta_ep, nsplit, nch_total, ch0, ch1 - ch0, sums);
    LAUNCHED(c);
    if (c->world > 1 && c->comm)
        NC(g_nccl.AllReduce(sums, sums, (size_t)2 * nch_total, ncclDouble, ncclSum, c->comm, c->st));
    k_sum2<<<1, 1024, 0, c->st>>>(sums, sums + nch_total, nch_total, c->d_energy);
    LAUNCHED(c);
    CU(cudaMemcpyAsync(c->pin->energy, c->d_energy, 2 * sizeof(double), cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    if (ms_out) CU(cudaEventElapsedTime(ms_out, c->ev0, c->ev1));
    return 0;
}
}  // namespace

extern "C" int gdy_energy(gdy_ctx *c, double *ekin_out, double *epot_out)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    int r;
    if ((r = need_particles(c))) return r;
    if ((r = energy_launch(c, nullptr))) return r;
    if (ekin_out) *ekin_out = c->pin->energy[0];
    if (epot_out) *epot_out = c->pin->energy[1];
    return GDY_OK;
}

// the same with the pair kernel bracketed by CUDA events (bench: roofline of the FP64-bound energy kernel)
extern "C" int gdy_energy_timed(gdy_ctx *c, double *ekin_out, double *epot_out, float *ms_out)
{
    if (!c || !ms_out) return GDY_ERR_INVALID;
    BIND(c);
    int r;
    if ((r = need_particles(c))) return r;
    if ((r = energy_launch(c, ms_out))) return r;
    if (ekin_out) *ekin_out = c->pin->energy[0];
    if (epot_out) *epot_out = c->pin->energy[1];
    return GDY_OK;
}

namespace {
int knn_launch(gdy_ctx *c, int k, double *dist_out, float *ms_out)
{
    if (k < 1 || k > KNN_MAX) return fail(c, GDY_ERR_INVALID, "k must be in [1, %d]", KNN_MAX);
    if (c->n <= (uint32_t)k) return fail(c, GDY_ERR_INVALID, "the %d-th neighbour needs more than %d particles (n = %u)", k, k, c->n);
    int r;
    if ((r = need_particles(c))) return r;
    double *d = c->s.part;   // scratch
    if (ms_out) CU(cudaEventRecord(c->ev0, c->st));
    k_knn_distance<<<cdiv(c->n, KNN_T), KNN_T, 0, c->st>>>(c->s, c->n, c->npad, k, d);
    LAUNCHED(c);
    if (ms_out) CU(cudaEventRecord(c->ev1, c->st));
    if (dist_out) CU(cudaMemcpyAsync(dist_out, d, (size_t)c->n * sizeof(double), cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    if (ms_out) CU(cudaEventElapsedTime(ms_out, c->ev0, c->ev1));
    return 0;
}
}  // namespace

extern "C" int gdy_neighbour_distance(gdy_ctx *c, int k, double *dist_out)
{
    if (!c) return GDY_ERR_INVALID;
    if (!dist_out) return fail(c, GDY_ERR_INVALID, "dist_out is NULL");
    BIND(c);
    return knn_launch(c, k, dist_out, nullptr);
}

extern "C" int gdy_neighbour_timed(gdy_ctx *c, int k, float *ms_out)
{
    if (!c || !ms_out) return GDY_ERR_INVALID;
    BIND(c);
    return knn_launch(c, k, nullptr, ms_out);
}

extern "C" uint64_t gdy_launch_count(const gdy_ctx *c) { return c ? c->launches : 0; }

extern "C" int gdy_exchange_stats(const gdy_ctx *c, uint64_t *peer_out, uint64_t *nccl_out)
{
    if (!c) return GDY_ERR_INVALID;
    if (peer_out) *peer_out = c->xmid_count;
    if (nccl_out) *nccl_out = c->nccl_count;
    return GDY_OK;
}

extern "C" int gdy_sync(gdy_ctx *c)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    CU(cudaStreamSynchronize(c->st));
    return check_abort(c);
}

// ---- persistent small-step launches ------------------------------------------------------------------------
namespace {

// Runs block steps inside ONE cooperative launch while the active set stays <= ps_nact_max (see
// gdy_persist.cuh).  Entry: ATIME = `itime` with its active list in State.move[0..nact).  Returns with
// the pinned PersistCtl filled in; State.move then holds the list of the first step NOT done.
// The second design needs its slices' t+dt and predicted set in shared memory; how much of the raw state stays
// resident is what is left of the 227 KiB.  Returns GDY_OK with p2_ready == false when the slices do not fit
// (very large N per GPU): those contexts keep to the launch-per-hook path for their small steps.
int persist2_prepare(gdy_ctx *c)
{
    if (c->p2_ready || c->p2_smem == 1) return 0;
    const uint32_t G = c->ps_grid;
    uint32_t jb = c->blk_begin * JB, je = c->blk_end * JB < c->n ? c->blk_end * JB : c->n;
    if (je < jb) je = jb;
    const uint32_t lf = cdiv(je - jb, G) + 1, ls = cdiv(c->n, G) + 1;
    int max_optin = 0;
    CU(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device));
    const size_t off = (sizeof(P2Fixed) + 15) & ~(size_t)15;
    const uint32_t lf_pad = round_up(lf, 2), ls_cap = round_up(ls, 2);
    const size_t need = off + (size_t)7 * lf_pad * 8 + (size_t)ls_cap * 8;
    if (need > (size_t)max_optin) { c->p2_smem = 1; return 0; }   // does not fit: marked, not an error
    uint32_t res = (uint32_t)(((size_t)max_optin - need) / (13 * 8));
    if (res > lf_pad) res = lf_pad;
    if (const char *e = getenv("GDY_P2_RESIDENT")) { const long r = atol(e); if (r >= 0 && (uint32_t)r < res) res = (uint32_t)r; }
    c->p2_off_dyn = (uint32_t)off; c->p2_lf_pad = lf_pad; c->p2_ls_cap = ls_cap; c->p2_res_cap = res;
    c->p2_smem = (uint32_t)(need + (size_t)13 * res * 8);
    // the attribute belongs to the FUNCTION, not to this context: always the device maximum, so that contexts with
    // different carve-ups (two logical ranks in one process) cannot shrink it under each other
    CU(cudaFuncSetAttribute((const void *)k_persist2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin));
    CU(cudaFuncSetAttribute((const void *)k_persist2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin));
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void *)k_persist2<false>, PS_T, c->p2_smem));
    if (per_sm < 1) { c->p2_smem = 1; return 0; }
    const size_t gp = c->ps_gpad;
    auto zalloc = [&](void **p, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(p, bytes);
        return e != cudaSuccess ? e : cudaMemsetAsync(*p, 0, bytes, c->st);
    };
    CU(zalloc((void **)&c->p2_ipub, (size_t)PS_NACT_CAP * 6 * sizeof(uint4)));
    CU(zalloc((void **)&c->p2_part, (size_t)PS_NACT_CAP * 6 * gp * sizeof(uint4)));
    CU(zalloc((void **)&c->p2_anext, (size_t)PS_NACT_CAP * sizeof(uint4)));
    CU(zalloc((void **)&c->p2_cmin, 2 * gp * sizeof(uint4)));
    CU(zalloc((void **)&c->p2_ccnt, 2 * gp * sizeof(uint2)));
    CU(zalloc((void **)&c->p2_clist, (size_t)2 * G * PS_NACT_CAP * sizeof(uint2)));
    CU(zalloc((void **)&c->p2_ecnt, gp * sizeof(uint2)));
    CU(zalloc((void **)&c->p2_abort, sizeof(uint32_t)));
    if (getenv("GDY_PERSIST_PROF")) CU(zalloc((void **)&c->p2_prof, P2_NPROF * sizeof(unsigned long long)));
    CU(cudaStreamSynchronize(c->st));
    c->p2_ready = true;
    return 0;
}

int run_persist(gdy_ctx *c, double itime, uint32_t nact, double t_stop, uint32_t budget, bool timed, float *ms_out)
{
    Persist2Args a{};
    a.s = c->s; a.n = c->n;
    a.jb = c->blk_begin * JB;
    a.je = c->blk_end * JB < c->n ? c->blk_end * JB : c->n;
    if (a.je < a.jb) a.je = a.jb;
    a.e2 = c->e2; a.eta = (double)c->eta; a.t_stop = t_stop;
    a.entry_itime = itime; a.entry_nact = nact;
    a.max_steps = budget < PS_STEPS_MAX ? budget : PS_STEPS_MAX;
    a.nact_max = c->ps_nact_max;
    a.seq = ++c->ps_seq;
    a.tag0 = c->ps_xseq;
    a.ipub = c->p2_ipub; a.part = c->p2_part; a.anext = c->p2_anext; a.cmin = c->p2_cmin; a.ccnt = c->p2_ccnt;
    a.clist = c->p2_clist; a.ecnt = c->p2_ecnt; a.gpad = c->ps_gpad; a.lcap = PS_NACT_CAP;
    a.off_dyn = c->p2_off_dyn; a.lf_pad = c->p2_lf_pad; a.res_cap = c->p2_res_cap; a.ls_cap = c->p2_ls_cap;
    a.ctl_dev = c->ps_ctl; a.ctl_host = c->ps_ctl_host; a.trace = c->ps_trace_host; a.prof = c->p2_prof;
    a.prof_nact_max = 0xffffffffu;
    if (const char *e = getenv("GDY_PERSIST_PROF_NACT")) { const long v = atol(e); if (v > 0) a.prof_nact_max = (uint32_t)v; }
    a.abort_word = c->p2_abort; a.timeout_ns = c->spin_timeout_ns;
    a.rank = c->rank; a.world = c->peers_ok ? c->world : 1;
    for (int r = 0; r < PS_MAX_WORLD; ++r) a.xslot[r] = c->xslot[r];
    void *args[] = {&a};
    const void *fn = c->mask_self ? (const void *)k_persist2<true> : (const void *)k_persist2<false>;
    {
        cudaError_t le = cudaLaunchCooperativeKernel(fn, dim3(c->ps_grid), dim3(PS_T), args, c->p2_smem, c->st);
        if (le != cudaSuccess)
        {
            int occ = -1;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, PS_T, c->p2_smem);
            return fail(c, GDY_ERR_CUDA, "cooperative launch of the persistent kernel: %s (grid %u x %d threads, %u B shared memory, "
                        "%d resident CTAs per SM x %d SMs)", cudaGetErrorString(le), c->ps_grid, (int)PS_T, c->p2_smem, occ, c->num_sms);
        }
    }
    ++c->launches; ++c->ps_launches;
    // CTA 0 writes the control block with `seq` last when it leaves: polled, not cudaStreamSynchronize
    { int r = wait_pinned(c, &c->ps_ctl_host->seq, a.seq, "the persistent kernel"); if (r) return r; }
    if (c->ps_ctl_host->epoch_end)
    {
        uint32_t ab = 0;
        CU(cudaStreamSynchronize(c->st));
        CU(cudaMemcpy(&ab, c->p2_abort, sizeof ab, cudaMemcpyDeviceToHost));
        return fail(c, GDY_ERR_STATE, "persistent kernel watchdog: rank %d gave up after %.0f ms in phase %u of CTA %u at block step %u "
                    "of the launch (a peer rank died, was never attached or runs a different schedule); the state is undefined",
                    c->rank, c->spin_timeout_ns * 1e-6, ab ? (ab - 1) >> 16 : 0u, ab ? (ab - 1) & 0xffffu : 0u, c->ps_ctl_host->steps);
    }
    // tags: one per step, one for the exit exchange, and one more so that the next launch's first step has the other
    // exchange-slot parity than this launch's last
    c->ps_xseq += c->ps_ctl_host->steps + 2;
    c->ps_steps_total += c->ps_ctl_host->steps;
    // the time of the launch's block steps by the kernel's own clock (%globaltimer stamps of CTA 0 in the trace)
    if (ms_out) *ms_out = (timed && c->ps_ctl_host->steps) ? c->ps_trace_host[c->ps_ctl_host->steps - 1].ns * 1e-6f : 0.f;
    return 0;
}

}  // namespace

extern "C" int gdy_set_persist(gdy_ctx *c, int nact_max)
{
    if (!c) return GDY_ERR_INVALID;
    if (nact_max > PS_NACT_CAP) return fail(c, GDY_ERR_INVALID, "nact_max %d > %d", nact_max, (int)PS_NACT_CAP);
    if (c->world > 1 && !c->peers_ok && nact_max != 0)
        return fail(c, GDY_ERR_STATE, "a sharded context needs gdy_peer_attach before the persistent path");
    c->ps_nact_max = nact_max < 0 ? PS_DEFAULT_NACT : (uint32_t)nact_max;
    return GDY_OK;
}

// Whether the persistent path can serve this context, answered without waiting for the first gdy_run (which is where
// persist2_prepare decides it): the same shared-memory sum as there.
extern "C" int gdy_persist_state(gdy_ctx *c, uint32_t *smem_need_out, uint32_t *smem_max_out)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    const uint32_t G = c->ps_grid ? c->ps_grid : (uint32_t)c->num_sms;
    uint32_t jb = c->blk_begin * JB, je = c->blk_end * JB < c->n ? c->blk_end * JB : c->n;
    if (je < jb) je = jb;
    const uint32_t lf = cdiv(je - jb, G) + 1, ls = cdiv(c->n, G) + 1;
    int max_optin = 0;
    CU(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device));
    const size_t off = (sizeof(P2Fixed) + 15) & ~(size_t)15;
    const size_t need = off + (size_t)7 * round_up(lf, 2) * 8 + (size_t)round_up(ls, 2) * 8;
    if (smem_need_out) *smem_need_out = (uint32_t)need;
    if (smem_max_out) *smem_max_out = (uint32_t)max_optin;
    if (c->ps_nact_max == 0) return 0;
    if (c->p2_smem == 1 || need > (size_t)max_optin) return 2;
    return 1;
}

// ---- peer memory for the fused exchange of the persistent path ------------------------------------------------
namespace {
struct PeerHandle   // GDY_PEER_HANDLE_BYTES
{
    uint64_t magic;
    int32_t pid, device, rank, world;
    uint64_t ptr;
    cudaIpcMemHandle_t ipc;
    uint8_t pad[GDY_PEER_HANDLE_BYTES - 8 - 16 - 8 - sizeof(cudaIpcMemHandle_t)];
};
static_assert(sizeof(PeerHandle) == GDY_PEER_HANDLE_BYTES, "peer handle size");
constexpr uint64_t PEER_MAGIC = 0x4744595f50454552ull;   // "GDY_PEER"
}  // namespace

extern "C" int gdy_peer_export(gdy_ctx *c, uint8_t *handle_out)
{
    if (!c || !handle_out) return GDY_ERR_INVALID;
    BIND(c);
    if (!c->xbuf)
    {
        CU(cudaMalloc(&c->xbuf, (XBUF_SLOTS + XMID_SLOTS) * sizeof(uint4)));
        CU(cudaMemset(c->xbuf, 0, (XBUF_SLOTS + XMID_SLOTS) * sizeof(uint4)));
    }
    PeerHandle h{};
    h.magic = PEER_MAGIC; h.pid = (int32_t)getpid(); h.device = c->device; h.rank = c->rank; h.world = c->world;
    h.ptr = (uint64_t)(uintptr_t)c->xbuf;
    CU(cudaIpcGetMemHandle(&h.ipc, c->xbuf));
    memcpy(handle_out, &h, sizeof h);
    return GDY_OK;
}

extern "C" int gdy_peer_attach(gdy_ctx *c, const uint8_t *handles)
{
    if (!c || !handles) return GDY_ERR_INVALID;
    BIND(c);
    if (c->world < 2) return GDY_OK;
    if (c->world > PS_MAX_WORLD) return fail(c, GDY_ERR_INVALID, "world %d > %d", c->world, (int)PS_MAX_WORLD);
    if (!c->xbuf) return fail(c, GDY_ERR_STATE, "gdy_peer_attach before gdy_peer_export");
    for (int r = 0; r < c->world; ++r)
    {
        PeerHandle h;
        memcpy(&h, handles + (size_t)r * GDY_PEER_HANDLE_BYTES, sizeof h);
        if (h.magic != PEER_MAGIC || h.rank != r || h.world != c->world)
            return fail(c, GDY_ERR_INVALID, "peer handle %d is not rank %d of %d", r, r, c->world);
        if (r == c->rank) { c->xslot[r] = c->xbuf; continue; }
        if (h.pid == (int32_t)getpid() && h.device == c->device)
            c->xslot[r] = (uint4 *)(uintptr_t)h.ptr;   // two logical ranks on one device (tests): the same address space
        else if (h.pid == (int32_t)getpid())
        {
            // same process (one thread per GPU): plain peer access to the other context's allocation
            int can = 0;
            CU(cudaDeviceCanAccessPeer(&can, c->device, h.device));
            if (!can) return fail(c, GDY_ERR_CUDA, "device %d cannot access device %d", c->device, h.device);
            cudaError_t e = cudaDeviceEnablePeerAccess(h.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                return fail(c, GDY_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d): %s", h.device, cudaGetErrorString(e));
            (void)cudaGetLastError();
            c->xslot[r] = (uint4 *)(uintptr_t)h.ptr;
        }
        else
        {
            void *p = nullptr;
            CU(cudaIpcOpenMemHandle(&p, h.ipc, cudaIpcMemLazyEnablePeerAccess));
            c->xopened[r] = p;
            c->xslot[r] = (uint4 *)p;
        }
    }
    c->peers_ok = true;
    // on four GPUs and more the launch path's fixed cost per block step (launch chain + exchange, ~80 us) outweighs the
    // persistent kernel's lower pair rate up to its buffer cap: N = 256k on 8 GPUs to t = 1/8 in 2.041 s with 2048 against
    // 2.075 s with 512 (profiles/r02_persist_8gpu.md); on one GPU 512 stays (blocks of 513..1024: 622 against 584 us)
    uint32_t def = c->world >= 4 ? (uint32_t)PS_NACT_CAP : PS_DEFAULT_NACT;
    if (const char *e = getenv("GDY_PERSIST_NACT")) def = (uint32_t)atoi(e);
    if (def > (uint32_t)PS_NACT_CAP) def = PS_NACT_CAP;
    c->ps_nact_max = def;
    return GDY_OK;
}

// ns spent per phase of the persistent kernel by CTA 0 (only with GDY_PERSIST_PROF=1 in the environment)
extern "C" int gdy_persist_profile(gdy_ctx *c, uint64_t *ns_out /* [144]: the first P2_NPROF = 16 are used */)
{
    if (!c || !ns_out) return GDY_ERR_INVALID;
    BIND(c);
    memset(ns_out, 0, 144 * sizeof(uint64_t));
    if (!c->p2_prof) return GDY_OK;
    CU(cudaMemcpyAsync(ns_out, c->p2_prof, P2_NPROF * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    return GDY_OK;
}

extern "C" int gdy_persist_stats(const gdy_ctx *c, uint64_t *launches_out, uint64_t *steps_out)
{
    if (!c) return GDY_ERR_INVALID;
    if (launches_out) *launches_out = c->ps_launches;
    if (steps_out) *steps_out = c->ps_steps_total;
    return GDY_OK;
}

// ---- fused driver ----------------------------------------------------------------------------------------
// The reference loop body (src/include/cpu/Hermite4CPU_integration.cpp:81-124) with everything on the
// device and ONE host readback per block step ({next block time, next nact}, 16 bytes of pinned
// memory): the scheduler of step k+1 is chained behind the corrector of step k.
extern "C" int gdy_run(gdy_ctx *c, double atime, double itime0, double t_stop, uint64_t max_steps,
                       gdy_run_stats *stats, uint32_t *trace_nact, double *trace_itime, float *trace_us,
                       uint64_t trace_cap)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    c->sched_cached = false;
    int r;
    if ((r = need_particles(c))) return r;
    State &s = c->s;
    const auto w0 = std::chrono::steady_clock::now();
    const bool timed = stats != nullptr;   // two event records per block step: force_ms feeds the log's GFLOPS
    uint64_t it = 0, nsteps = 0, inter = 0;
    double force_ms = 0.0;
    double ITIME = itime0, ATIME = atime;

    if (ITIME < t_stop && !(max_steps && it >= max_steps))
    {
        if ((r = launch_sched(c, ATIME, false))) return r;
        CU(cudaStreamSynchronize(c->st));
    }
    auto tick = std::chrono::steady_clock::now();
    while (ITIME < t_stop)
    {
        if (max_steps && it >= max_steps) break;
        const uint32_t nact = c->pin->sched.nact;
        if (c->ps_nact_max && !c->p2_ready && c->p2_smem != 1 && (r = persist2_prepare(c))) return r;
        if (c->ps_nact_max && nact && nact <= c->ps_nact_max && c->p2_ready)
        {
            // small active set: stay on the device until a big block (or t_stop / the step budget) comes up
            const uint64_t left = max_steps ? max_steps - it : PS_STEPS_MAX;
            const uint32_t budget = (uint32_t)(left < PS_STEPS_MAX ? left : PS_STEPS_MAX);
            float ms = 0.f;
            if ((r = run_persist(c, ATIME, nact, t_stop, budget, timed, &ms))) return r;
            const PersistCtl pc = *c->ps_ctl_host;
            if (pc.steps == 0) return fail(c, GDY_ERR_STATE, "persistent kernel made no progress");
            force_ms += ms;
            if (it < trace_cap && (trace_nact || trace_itime || trace_us))
            {
                const uint64_t k = pc.steps < trace_cap - it ? pc.steps : trace_cap - it;
                const auto now = std::chrono::steady_clock::now();
                const float us = std::chrono::duration<float, std::micro>(now - tick).count();
                tick = now;
                // per-step time = the kernel's own stamps; what the launch cost around them goes to its first step
                const float in_kernel = c->ps_trace_host[k - 1].ns * 1e-3f;
                for (uint64_t q = 0; q < k; ++q)
                {
                    if (trace_nact) trace_nact[it + q] = c->ps_trace_host[q].nact;
                    if (trace_itime) trace_itime[it + q] = c->ps_trace_host[q].itime;
                    if (trace_us)
                        trace_us[it + q] = q ? (c->ps_trace_host[q].ns - c->ps_trace_host[q - 1].ns) * 1e-3f
                                             : c->ps_trace_host[0].ns * 1e-3f + (k == pc.steps && us > in_kernel ? us - in_kernel : 0.f);
                }
            }
            else tick = std::chrono::steady_clock::now();
            it += pc.steps; nsteps += pc.sum_nact; inter += pc.sum_nact * (uint64_t)c->n;
            ITIME = pc.last_itime; ATIME = pc.atime;
            c->last_nact = pc.last_nact;
            c->pin->sched.nact = pc.nact_next;
            c->pin->sched.itime = pc.atime;
            c->ps_last = true;
            if (!pc.move_is_next && ITIME < t_stop && !(max_steps && it >= max_steps))
            {
                // stopped on the per-launch step cap: State.move still holds the last completed step's list
                if ((r = launch_sched(c, ATIME, false))) return r;
                CU(cudaStreamSynchronize(c->st));
            }
            continue;
        }
        c->ps_last = false;
        ITIME = ATIME;
        c->last_nact = nact;
        const bool last = !(ITIME < t_stop) || (max_steps && it + 1 >= max_steps);
        // nact == 0 happens in the reference too: on a resumed run init_dt returns the smallest STEP as
        // the first block time although every particle sits at the snapshot time, so the first pass
        // through the loop moves nobody and only finds the real next block time.  Same here.
        const double *buf = nullptr; uint32_t ld = 0;
        if (nact)
        {
            if ((r = launch_predict(c, ITIME, nact))) return r;
            XRecv xr;
            if ((r = force_eval(c, nact, s.move, &buf, &ld, timed, &xr))) return r;
            k_correct<<<cdiv(nact, 128), 128, 0, c->st>>>(s, nact, ITIME, (double)c->eta, buf, ld, xr);
            LAUNCHED(c);
        }
        if ((r = launch_min_time(c))) return r;
        if (last)
        {
            ++c->seq;
            k_min_final<<<1, 256, 0, c->st>>>(c->d_block_min, c->n_min_blocks, c->d_atime, &c->pin->atime, &c->pin->atime_seq, c->seq);
            LAUNCHED(c);
        }
        else if ((r = launch_sched(c, 0.0, true))) return r;
        if (last) CU(cudaStreamSynchronize(c->st));
        else if ((r = wait_sched(c))) return r;
        if ((r = check_abort(c))) return r;
        ATIME = last ? c->pin->atime : c->pin->sched.itime;
        if (timed && nact)
        {
            float ms = 0.f;
            cudaError_t ee = cudaEventElapsedTime(&ms, c->ev0, c->ev1);
            if (ee == cudaErrorNotReady)   // the slot can be ahead of the event's own completion record
            {
                (void)cudaGetLastError();
                CU(cudaEventSynchronize(c->ev1));
                ee = cudaEventElapsedTime(&ms, c->ev0, c->ev1);
            }
            CU(ee);
            force_ms += ms;
        }
        if (it < trace_cap)
        {
            if (trace_nact) trace_nact[it] = nact;
            if (trace_itime) trace_itime[it] = ITIME;
            if (trace_us)
            {
                const auto now = std::chrono::steady_clock::now();
                trace_us[it] = std::chrono::duration<float, std::micro>(now - tick).count();
                tick = now;
            }
        }
        inter += (uint64_t)nact * (uint64_t)c->n;
        nsteps += nact;
        ++it;
    }
    if (c->ps_last)
    {
        // h_p as the reference leaves it: the persistent kernel predicts on the fly and stores only the
        // active particles' predictions; the others are recomputed here (same arithmetic, same bits)
        k_predict_rest<<<cdiv(c->n, 256), 256, 0, c->st>>>(s, c->n, ITIME);
        LAUNCHED(c);
        c->have_pred = true;
        c->ps_last = false;
    }
    if (stats)
    {
        stats->iterations = it; stats->nsteps = nsteps; stats->interactions = inter;
        stats->itime = ITIME; stats->atime = ATIME; stats->force_ms = force_ms;
        stats->wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count();
        stats->last_nact = it ? c->last_nact : 0;
        stats->reserved = 0;
    }
    return GDY_OK;
}

// ---- bench / test entry points -----------------------------------------------------------------------------
extern "C" int gdy_force_sweep(gdy_ctx *c, uint32_t nact, const uint32_t *move, double *f6_out, float *ms_out)
{
    if (!c) return GDY_ERR_INVALID;
    if (nact == 0 || nact > c->n || !move) return fail(c, GDY_ERR_INVALID, "bad active list (nact %u, n %u)", nact, c->n);
    BIND(c);
    c->sched_cached = false;
    if (!c->have_pred) return fail(c, GDY_ERR_STATE, "force_sweep before predicted_pos_vel/init_acc_jrk");
    for (uint32_t k = 0; k < nact; ++k)
        if (move[k] >= c->n) return fail(c, GDY_ERR_INVALID, "move[%u] = %u out of range", k, move[k]);
    CU(cudaMemcpyAsync(c->d_ilist, move, (size_t)nact * sizeof(uint32_t), cudaMemcpyHostToDevice, c->st));
    const double *buf; uint32_t ld;
    int r;
    if ((r = force_eval(c, nact, c->d_ilist, &buf, &ld, true))) return r;
    if (f6_out)
    {
        std::vector<double> tmp((size_t)6 * ld);
        CU(cudaMemcpyAsync(tmp.data(), buf, tmp.size() * sizeof(double), cudaMemcpyDeviceToHost, c->st));
        CU(cudaStreamSynchronize(c->st));
        for (uint32_t k = 0; k < nact; ++k)
            for (int q = 0; q < 6; ++q) f6_out[(size_t)k * 6 + q] = tmp[(size_t)q * ld + k];
    }
    else CU(cudaStreamSynchronize(c->st));
    if (ms_out) CU(cudaEventElapsedTime(ms_out, c->ev0, c->ev1));
    return GDY_OK;
}

extern "C" int gdy_force_sweep_full(gdy_ctx *c, int reps, float *ms_out)
{
    if (!c) return GDY_ERR_INVALID;
    if (reps < 1) return fail(c, GDY_ERR_INVALID, "reps must be >= 1");
    BIND(c);
    c->sched_cached = false;
    if (!c->have_pred) return fail(c, GDY_ERR_STATE, "force_sweep_full before predicted_pos_vel/init_acc_jrk");
    // the context's own event pair (force_eval leaves it alone when it is not asked to time): nothing to leak on an
    // error return
    CU(cudaEventRecord(c->ev0, c->st));
    const double *buf; uint32_t ld;
    for (int i = 0; i < reps; ++i)
    {
        int r = force_eval(c, c->n, nullptr, &buf, &ld, false);
        if (r) return r;
    }
    CU(cudaEventRecord(c->ev1, c->st));
    CU(cudaStreamSynchronize(c->st));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (ms_out) *ms_out = ms / reps;
    return GDY_OK;
}

// ---- roofline denominators (measurement infrastructure) ------------------------------------------------------
extern "C" int gdy_measure_fp64_peak(gdy_ctx *c, double *tflops_out, double *sm_mhz_out)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    const int iters = 4096;                            // x 64 DFMA per thread
    double *out = c->s.part;                           // scratch (>= 8 Mi doubles)
    using Fn = void (*)(double *, int, double, double);
    const Fn fns[5] = {k_dfma_peak<4>, k_dfma_peak<8>, k_dfma_peak<16>, k_dfma_peak_acc<8>, k_dfma_peak_acc<16>};
    double best_tf = 0.0;
    for (int f = 0; f < 5; ++f)
        for (int per_sm = 1; per_sm <= 8; per_sm *= 2)  // 256..2048 resident threads per SM
        {
            const int blocks = c->num_sms * per_sm;
            float best = 1e30f;
            for (int rep = 0; rep < 4; ++rep)
            {
                CU(cudaEventRecord(c->ev0, c->st));
                fns[f]<<<blocks, 256, 0, c->st>>>(out, iters, 1.0000001, 1e-9);
                LAUNCHED(c);
                CU(cudaEventRecord(c->ev1, c->st));
                CU(cudaStreamSynchronize(c->st));
                float ms;
                CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
                if (rep >= 1 && ms < best) best = ms;
            }
            const double tf = 2.0 * 64.0 * iters * 256.0 * blocks / (best * 1e-3) / 1e12;
            if (tf > best_tf) best_tf = tf;
        }
    if (tflops_out) *tflops_out = best_tf;
    // the SM clock this rate corresponds to if the pipe retires 64 DFMA lanes per SM per cycle
    if (sm_mhz_out) *sm_mhz_out = best_tf * 1e12 / (2.0 * 64.0 * c->num_sms) / 1e6;
    return GDY_OK;
}

extern "C" int gdy_measure_copy_bw(gdy_ctx *c, double *gbs_out)
{
    if (!c) return GDY_ERR_INVALID;
    BIND(c);
    const size_t bytes = (size_t)1 << 30;
    char *a, *b;
    CU(cudaMalloc(&a, bytes)); CU(cudaMalloc(&b, bytes));
    CU(cudaMemsetAsync(a, 1, bytes, c->st));
    float best = 1e30f;
    for (int rep = 0; rep < 6; ++rep)
    {
        CU(cudaEventRecord(c->ev0, c->st));
        CU(cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice, c->st));
        CU(cudaEventRecord(c->ev1, c->st));
        CU(cudaStreamSynchronize(c->st));
        float ms;
        CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        if (rep >= 1 && ms < best) best = ms;
    }
    cudaFree(a); cudaFree(b);
    if (gbs_out) *gbs_out = 2.0 * bytes / (best * 1e-3) / 1e9;
    return GDY_OK;
}

#include <mutex>

// ---- total energy for callers that must not allocate ---------------------------------------------------------
// gdy_energy_noalloc(): the same chunk / j-split / fold decomposition as energy_launch() above, but without a
// context.  Who needs it: a host program that gave the whole device to its own memory pool at start-up (a
// cudaMalloc from a library would fail, or eat into a budget the program thinks it owns); an analysis hook that runs
// between block steps and may not stall the GPU, which cudaMalloc/cudaFree do because they synchronise the device;
// a plug-in host whose rules forbid driver allocations after initialisation.  Such a caller already holds its
// particles in device memory; it passes them (and a two-double result slot) in, and the per-CTA partial sums, which
// gdy_energy keeps in the context's cudaMalloc'ed partial-force buffer, live in the fixed array below instead.
//
// g_noalloc_scratch is part of the library image: every GPU that loads libgravidy_b200.so carries these 128 KiB until
// the process exits, used or not.  All calls on a device share it and it keeps whatever the previous call left.
// Nothing in it has to start at zero: every slot a call reads (cta_ek[chunk], cta_ep[split][chunk], sums[2][chunk])
// is written earlier in the same call by exactly one thread, so stale contents are never seen.  Two calls that are in
// flight on the same GPU at the same time (different streams, or different host threads) write the same slots and race.
//
// In host code the identifier g_noalloc_scratch is only a handle for the runtime, not a pointer the GPU can follow,
// so the host asks for the device address with cudaGetSymbolAddress, once per device and with that device current,
// and hands that to the kernels.  This happens inside the first call on each device.
constexpr uint32_t NOALLOC_MAX_N = 1u << 20;          // GDY_NOALLOC_MAX_N in the header
constexpr size_t NOALLOC_SCRATCH_DOUBLES = 16384;     // need(n) = (nsplit + 3) * ceil(n / 512) <= 8192 for n <= 2^20
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SCRATCH_DOUBLES];

namespace {

// k_energy with the reference's host layout as input: r4 = n x {x, y, z, m}, v4 = n x {vx, vy, vz, _}, no padded
// copies.  Indices past n are staged as the padding particle of the context path (mass 0 at PAD_POS), masses are
// narrowed to float like gdy_upload_particles does, and the loop body, tile order and reduction are k_energy's.
__global__ void __launch_bounds__(EN_T)
k_energy_aos(const double *__restrict__ r4, const double *__restrict__ v4, uint32_t n, uint32_t tiles_per_split,
             uint32_t ntiles, double *__restrict__ cta_ek, double *__restrict__ cta_ep, uint32_t nch_total)
{
    __shared__ __align__(16) double tile[2][4][EN_TILE];
    __shared__ double red[8];
    const int tid = threadIdx.x;
    const uint32_t chunk = blockIdx.x;
    const uint32_t i0 = chunk * EN_CHUNK + tid, i1 = i0 + EN_T;
    const uint32_t c0 = i0 < n ? i0 : n - 1, c1 = i1 < n ? i1 : n - 1;
    const double x0 = r4[4 * (size_t)c0], y0 = r4[4 * (size_t)c0 + 1], z0 = r4[4 * (size_t)c0 + 2];
    const double x1 = r4[4 * (size_t)c1], y1 = r4[4 * (size_t)c1 + 1], z1 = r4[4 * (size_t)c1 + 2];
    const uint32_t t0 = blockIdx.y * tiles_per_split, t1 = min(t0 + tiles_per_split, ntiles);
    double p0 = 0.0, p1 = 0.0;
    auto stage = [&](uint32_t t, int b) {
        const uint32_t j = t * EN_TILE + tid;
        const bool real = j < n;
        const double *q = r4 + 4 * (size_t)(real ? j : 0);
        tile[b][0][tid] = real ? q[0] : PAD_POS;
        tile[b][1][tid] = real ? q[1] : PAD_POS;
        tile[b][2][tid] = real ? q[2] : PAD_POS;
        tile[b][3][tid] = real ? (double)(float)q[3] : 0.0;
    };
    if (t0 < t1) stage(t0, 0);
    __syncthreads();
    for (uint32_t t = t0; t < t1; ++t)
    {
        const int b = (t - t0) & 1;
        if (t + 1 < t1) stage(t + 1, b ^ 1);
#pragma unroll 4
        for (int j = 0; j < EN_TILE; j += 2)
        {
            const double2 X = *reinterpret_cast<const double2 *>(&tile[b][0][j]);
            const double2 Y = *reinterpret_cast<const double2 *>(&tile[b][1][j]);
            const double2 Z = *reinterpret_cast<const double2 *>(&tile[b][2][j]);
            const double2 M = *reinterpret_cast<const double2 *>(&tile[b][3][j]);
#define GDY_EN_PAIR(xi, yi, zi, xj, yj, zj, mj, pot) \
            { \
                const double dx = xj - xi, dy = yj - yi, dz = zj - zi; \
                const double r2 = fma(dz, dz, fma(dy, dy, dx * dx)); \
                double y = rsqrt_f64(r2); \
                y = (r2 > 0.0) ? y : 0.0; \
                pot = fma(mj, y, pot); \
            }
            GDY_EN_PAIR(x0, y0, z0, X.x, Y.x, Z.x, M.x, p0)
            GDY_EN_PAIR(x1, y1, z1, X.x, Y.x, Z.x, M.x, p1)
            GDY_EN_PAIR(x0, y0, z0, X.y, Y.y, Z.y, M.y, p0)
            GDY_EN_PAIR(x1, y1, z1, X.y, Y.y, Z.y, M.y, p1)
#undef GDY_EN_PAIR
        }
        __syncthreads();
    }
    const double m0 = i0 < n ? (double)(float)r4[4 * (size_t)i0 + 3] : 0.0;
    const double m1 = i1 < n ? (double)(float)r4[4 * (size_t)i1 + 3] : 0.0;
    double ep = 0.0, ek = 0.0;
    if (i0 < n) ep += -0.5 * m0 * p0;
    if (i1 < n) ep += -0.5 * m1 * p1;
    ep = block_sum_256(ep, red);
    if (tid == 0) cta_ep[(size_t)blockIdx.y * nch_total + chunk] = ep;
    if (blockIdx.y == 0)
    {
        if (i0 < n) { const double *w = v4 + 4 * (size_t)i0; const double vx = w[0], vy = w[1], vz = w[2]; ek += 0.5 * m0 * (vx * vx + vy * vy + vz * vz); }
        if (i1 < n) { const double *w = v4 + 4 * (size_t)i1; const double vx = w[0], vy = w[1], vz = w[2]; ek += 0.5 * m1 * (vx * vx + vy * vy + vz * vz); }
        ek = block_sum_256(ek, red);
        if (tid == 0) cta_ek[chunk] = ek;
    }
}

// device address of g_noalloc_scratch, looked up once per device (the first call on that device does it)
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
double *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

int noalloc_scratch(double **out)
{
    gdy_ctx *c = nullptr;   // errors go to the slot gdy_last_error(NULL) reads
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(c, GDY_ERR_INVALID, "gdy_energy_noalloc: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev])
    {
        int major = 0;
        CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major < 10) return fail(c, GDY_ERR_NO_DEVICE, "gdy_energy_noalloc: device %d is not sm_100 (there is no CPU path)", dev);
        double *p = nullptr;
        CU(cudaGetSymbolAddress((void **)&p, g_noalloc_scratch));
        if (!p) return fail(c, GDY_ERR_CUDA, "gdy_energy_noalloc: cudaGetSymbolAddress gave a null address on device %d", dev);
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

}  // namespace

extern "C" int gdy_energy_noalloc(const double *d_r4, const double *d_v4, uint32_t n, double *d_energy_out, void *stream)
{
    gdy_ctx *c = nullptr;
    if (!d_r4 || !d_v4 || !d_energy_out) return fail(c, GDY_ERR_INVALID, "gdy_energy_noalloc: null device pointer");
    if (n == 0) return fail(c, GDY_ERR_INVALID, "gdy_energy_noalloc: n = 0");
    if (n > NOALLOC_MAX_N)
        return fail(c, GDY_ERR_INVALID, "gdy_energy_noalloc: n = %u is above GDY_NOALLOC_MAX_N = %u, the size the library's fixed "
                    "scratch array is laid out for; use gdy_energy on a context for larger sets", n, NOALLOC_MAX_N);
    double *scratch = nullptr;
    int r;
    if ((r = noalloc_scratch(&scratch))) return r;
    cudaStream_t st = (cudaStream_t)stream;
    // the decomposition of energy_launch() for one GPU: a function of n alone
    const uint32_t nch_total = cdiv(n, EN_CHUNK), ntiles = round_up(n, PAD) / EN_TILE;
    uint32_t nsplit = cdiv(8u * 148u, nch_total);
    if (nsplit > ntiles / 4) nsplit = ntiles / 4 ? ntiles / 4 : 1;
    if (nsplit < 1) nsplit = 1;
    const uint32_t per = cdiv(ntiles, nsplit);
    nsplit = cdiv(ntiles, per);
    if ((size_t)(nsplit + 3) * nch_total > NOALLOC_SCRATCH_DOUBLES)
        return fail(c, GDY_ERR_INVALID, "gdy_energy_noalloc: n = %u needs %zu scratch doubles, the fixed array holds %zu",
                    n, (size_t)(nsplit + 3) * nch_total, NOALLOC_SCRATCH_DOUBLES);
    double *cta_ek = scratch, *cta_ep = scratch + nch_total, *sums = cta_ep + (size_t)nsplit * nch_total;
    k_energy_aos<<<dim3(nch_total, nsplit), EN_T, 0, st>>>(d_r4, d_v4, n, per, ntiles, cta_ek, cta_ep, nch_total);
    CU(cudaGetLastError());
    k_energy_fold<<<cdiv(nch_total, 256), 256, 0, st>>>(cta_ek, cta_ep, nsplit, nch_total, 0, nch_total, sums);
    CU(cudaGetLastError());
    k_sum2<<<1, 1024, 0, st>>>(sums, sums + nch_total, nch_total, d_energy_out);
    CU(cudaGetLastError());
    return GDY_OK;
}
