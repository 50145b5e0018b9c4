/* gravidy_b200.h — C ABI of the B200-native Hermite4 hot path.
 *
 * One `gdy_ctx` = one GPU (one j-shard).  It owns the device-resident particle state, streams and
 * (when world > 1) an NCCL communicator.  The entry points are the hooks the reference's
 * integration loop calls (reference: src/include/cpu/Hermite4CPU_integration.cpp:81-124, identical
 * in src/include/gpu/Hermite4GPU_integration.cpp:81-124); each one replaces the Hermite4GPU
 * member cited beside it.  Plain pointers and sizes only; no C++/torch types cross this boundary.
 *
 * Conventions
 *   - every call returns 0 on success, a negative gdy_status on error; gdy_last_error(ctx) gives
 *     the message.  (The reference's convention is print + exit(EXIT_FAILURE), reference:
 *     src/include/gpu/Hermite4GPU.cuh:43-50; the host shim in gravidy_b200/host keeps it.)
 *   - host pointers are borrowed for the duration of the call.
 *   - one caller thread per context (like the reference); calls are stream-ordered; only the calls
 *     that return a value to the host synchronise.
 *   - array shapes follow the reference's host arrays: r4 = N x {x,y,z,m} (double4 h_r), v4 = N x
 *     {x,y,z,_} (double4 h_v), f6 = N x {a[3],a1[3]} (Forces), a3x3 = N x {x,y,z} (double3).
 *   - there is NO CPU fallback: without a CUDA device gdy_create fails with GDY_ERR_NO_DEVICE.
 */
#ifndef GRAVIDY_B200_H
#define GRAVIDY_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gdy_ctx gdy_ctx;

typedef enum gdy_status
{
    GDY_OK = 0,
    GDY_ERR_INVALID = -1,   /* bad argument */
    GDY_ERR_NO_DEVICE = -2, /* no usable CUDA device: there is no CPU path */
    GDY_ERR_CUDA = -3,      /* CUDA runtime error, see gdy_last_error */
    GDY_ERR_NCCL = -4,      /* NCCL error or libnccl not loadable */
    GDY_ERR_STATE = -5      /* call sequence error (e.g. no particles uploaded) */
} gdy_status;

#define GDY_NCCL_UID_BYTES 128
#define GDY_MAX_N 67108863u /* 2^26 - 1 */

/* Version of this ABI (bumped on any change of a signature or of what an existing call means).
 *   3: gdy_energy_noalloc added; n limited to GDY_MAX_N; and the number now also covers what changed under version 2
 *      without a bump: gdy_upload_particles and gdy_energy are COLLECTIVE on sharded contexts that have a communicator
 *      (every rank must call them; a rank-0-only call blocks in NCCL), masses are stored as (double)(float)m like the
 *      reference's Predictor.m, and the host program's checkpoint format is GDYCKPT2.
 *   2: sharded contexts, persistent kernel, peer exchange.   1: round-1 hooks. */
int gdy_abi_version(void);

/* Number of usable sm_100 devices (0 if none or no driver): replaces cudaGetDeviceCount in
 * Hermite4GPU::Hermite4GPU (reference: src/include/gpu/Hermite4GPU.cu:50-60). */
int gdy_device_count(void);

/* ---- lifetime ------------------------------------------------------------------------------
 * Replaces Hermite4GPU::Hermite4GPU / alloc_arrays_device / ~Hermite4GPU
 * (reference: src/include/gpu/Hermite4GPU.cu:43-194).
 *   n       particle count, 1 <= n <= GDY_MAX_N (2^26 - 1: the index range the warp force kernel is built for;
 *           larger n is refused with GDY_ERR_INVALID rather than computed wrongly)
 *   e2      softening squared, as NbodySystem computes it (float product widened,
 *           reference: src/include/NbodySystem.cpp:50)
 *   eta     accuracy parameter, FLOAT as in the reference (src/include/NbodySystem.hpp:67)
 *   device  CUDA device ordinal for this context
 */
int gdy_create(gdy_ctx **out, uint32_t n, double e2, float eta, int device);

/* Sharded variant: this context is rank `rank` of `world` j-shards (contiguous equal slices,
 * reference partition: src/include/gpu/Hermite4GPU.cu:81-97).  State is replicated; partial forces
 * are summed with ncclAllReduce (replaces the host-side sum, reference: Hermite4GPU.cu:512-530).
 * nccl_uid: GDY_NCCL_UID_BYTES bytes from gdy_nccl_unique_id() on rank 0, distributed by the
 * caller (MPI, torch.distributed, a file, ...).  world == 1 ignores it (may be NULL).
 * Collective: returns when all `world` ranks have called it; a rank that waits longer than
 * GDY_NCCL_INIT_TIMEOUT_MS (environment, default 120000) for the others gets GDY_ERR_NCCL instead of
 * hanging (the reference reports and exits: src/include/gpu/Hermite4GPU.cuh:43-50).
 * After that timeout error the PROCESS MUST EXIT (the reference's convention): the helper thread that runs
 * ncclCommInitRank cannot be cancelled and stays blocked inside NCCL, bound to the device; if the missing rank turns
 * up later it completes a communicator nobody owns, and a second gdy_create_sharded with the same id would race with
 * it.  Do not retry in the same process. */
int gdy_create_sharded(gdy_ctx **out, uint32_t n, double e2, float eta, int device, int rank,
                       int world, const uint8_t *nccl_uid);
/* world > 1 with nccl_uid == NULL makes a LOGICAL shard: no communicator, force results are this
 * shard's partial sums and the caller adds them (single-GPU tests of the sharded path). */

/* Pure host arithmetic (no GPU needed): the particle range [begin, end) whose j-contributions shard
 * `rank` of `world` computes.  Contiguous, aligned to 64-particle blocks, covering [0, n). */
int gdy_shard_range(uint32_t n, int rank, int world, uint32_t *begin_out, uint32_t *end_out);
int gdy_nccl_unique_id(uint8_t *uid_out /* GDY_NCCL_UID_BYTES */);
int gdy_destroy(gdy_ctx *ctx);
const char *gdy_last_error(const gdy_ctx *ctx); /* ctx may be NULL: last create() error */

/* ---- state transfer ---------------------------------------------------------------------------
 * Replaces the per-step cudaMemcpy of the whole particle set (reference: Hermite4GPU.cu:226-265,
 * 338-346): state goes up once and comes down only at output points. */
/* r4 = n x {x, y, z, m}, v4 = n x {vx, vy, vz, unused}.  The mass is narrowed to float and widened again on the device,
 * (double)(float)m, because that is what the reference computes with (its reader and Predictor.m are float:
 * src/include/NbodySystem.cpp:174, src/include/common.hpp:216); masses that come from the reference's own reader are
 * unchanged by it.  Collective on sharded contexts with a communicator (every rank passes the same arrays, copies only
 * its 1/world of them over PCIe and the chunks are all-gathered over NVLink). */
int gdy_upload_particles(gdy_ctx *ctx, const double *r4, const double *v4);
int gdy_upload_times(gdy_ctx *ctx, const double *t, const double *dt); /* after host init_dt */
int gdy_upload_forces(gdy_ctx *ctx, const double *f6);                 /* tests / resume     */
int gdy_download_forces(gdy_ctx *ctx, double *f6);
/* Any pointer may be NULL.  old6 = Forces h_old, a2/a3 = double3 h_a2/h_a3, p7 = N x
 * {r[3],v[3],m} (Predictor), move = N x uint32 (entries >= nact are 0xFFFFFFFF as in the
 * reference, src/include/Hermite4.cpp:69). */
int gdy_download_state(gdy_ctx *ctx, double *r4, double *v4, double *f6, double *old6, double *a2,
                       double *a3, double *t, double *dt, double *p7, uint32_t *move);
/* Test hook: overwrite a2/a3/old on the device (NULL = keep). */
int gdy_upload_derivs(gdy_ctx *ctx, const double *old6, const double *a2, const double *a3);

/* ---- the hooks ---------------------------------------------------------------------------------- */
/* init_acc_jrk(): all-pairs acc+jerk at the uploaded r,v.  Replaces Hermite4GPU::init_acc_jrk +
 * k_init_acc_jrk (reference: Hermite4GPU.cu:333-375, Hermite4GPU_kernels.cu:47-87). */
int gdy_init_acc_jrk(gdy_ctx *ctx);

/* init_dt(): Aarseth initial steps on the HOST with libm log/pow exactly as the reference
 * (src/include/Hermite4.cpp:108-140) is done by the host shim; this convenience does
 * download_forces + that arithmetic + upload_times and returns the first block time. */
int gdy_init_dt(gdy_ctx *ctx, float eta_s, double itime, double *atime_out);

/* find_particles_to_move(ITIME): device compaction in ascending index order
 * (reference: src/include/Hermite4.cpp:63-84).  When ITIME is the time the preceding gdy_next_integration_time
 * returned and no call changed t or dt in between (the reference's loop), the answer is already there: that call ran
 * the scheduler for its result into a second list buffer, and this one only swaps the buffers — one host wait per
 * block step of the hook loop instead of two (11 us per block step, profiles/r02_hook_loop.md). */
int gdy_find_particles_to_move(gdy_ctx *ctx, double itime, uint32_t *nact_out);

/* save_old_acc_jrk(nact) (reference: src/include/Hermite4.cpp:145-153). */
int gdy_save_old_acc_jrk(gdy_ctx *ctx, uint32_t nact);

/* predicted_pos_vel(ITIME): replaces Hermite4GPU::predicted_pos_vel + k_prediction (reference:
 * Hermite4GPU.cu:201-268, Hermite4GPU_kernels.cu:127-158) with the CPU back end's expression order
 * (src/include/cpu/Hermite4CPU.cpp:340-365). */
int gdy_predicted_pos_vel(gdy_ctx *ctx, double itime);

/* update_acc_jrk(nact): replaces Hermite4GPU::update_acc_jrk + k_update + k_reduce + the host sum
 * (reference: Hermite4GPU.cu:385-535, Hermite4GPU_kernels.cu:165-248).  With world > 1 includes the
 * NCCL allreduce. */
int gdy_update_acc_jrk(gdy_ctx *ctx, uint32_t nact);

/* correction_pos_vel(ITIME, nact) incl. get_timestep_normal + normalize_dt (reference:
 * src/include/cpu/Hermite4CPU.cpp:370-411, src/include/utils/NbodyUtils.cpp:337-424). */
int gdy_correction_pos_vel(gdy_ctx *ctx, double itime, uint32_t nact);

/* next_integration_time(ATIME) (reference: src/include/Hermite4.cpp:90-102).  The active list of the current step
 * (the `move` array of gdy_download_state) stays as the last gdy_find_particles_to_move left it. */
int gdy_next_integration_time(gdy_ctx *ctx, double *atime_out);

/* get_energy_gpu() + k_energy (reference: Hermite4GPU.cu:540-592, Hermite4GPU_kernels.cu:250-288;
 * CPU: src/include/utils/NbodyUtils.cpp:427-529).  Unsoftened potential.  Sharded contexts with a communicator
 * split the i-particles in chunks of 512 over the ranks (collective call) and all get the same bits as a one-GPU
 * run; a sharded context WITHOUT a communicator returns the sum over its own chunks only. */
int gdy_energy(gdy_ctx *ctx, double *ekin_out, double *epot_out);
/* The same with the pair kernel bracketed by CUDA events (bench: FP64 roofline of k_energy). */
int gdy_energy_timed(gdy_ctx *ctx, double *ekin_out, double *epot_out, float *ms_out);

/* Total energy WITHOUT a context and without any allocation.  This entry point has no counterpart in the
 * reference (there get_energy_gpu works on the arrays alloc_arrays_device cudaMalloc'ed); it exists for callers that
 * cannot let a library allocate device memory: a program that handed the whole GPU to its own memory pool when it
 * started (a later cudaMalloc fails or breaks its accounting), an analysis hook between block steps that must not stall
 * the GPU (cudaMalloc and cudaFree synchronise the device), a plug-in host that forbids driver allocations after
 * start-up.  gdy_create / gdy_energy need cudaMalloc and pinned host memory, so those callers cannot use them.
 *
 *   d_r4          DEVICE pointer, n x {x, y, z, m} doubles (the reference's double4 h_r layout); masses are narrowed
 *                 to float precision like gdy_upload_particles does
 *   d_v4          DEVICE pointer, n x {vx, vy, vz, _} doubles (double4 h_v)
 *   n             1 <= n <= GDY_NOALLOC_MAX_N; anything larger is refused with GDY_ERR_INVALID and a message
 *   d_energy_out  DEVICE pointer, 2 doubles: kinetic, potential (unsoftened), written by the last kernel
 *   stream        a cudaStream_t (NULL = the default stream).  The call only enqueues three kernels; the result is
 *                 there once the stream has passed them.  Nothing is copied to the host and nothing is synchronised.
 * All four pointers belong to the device that is CURRENT on the calling thread.
 *
 * The per-CTA partial sums go to a fixed 128 KiB array that is part of the library image: every GPU that loads this
 * library holds it until the process exits, whether or not this function is ever called, and GDY_NOALLOC_MAX_N is
 * what fits in it.  The array is shared by all calls on a device and keeps the previous call's contents; nothing in
 * it needs clearing (every slot a call reads is written earlier in that call), but two calls that overlap on one GPU
 * (two streams, two threads) overwrite each other's partial sums: serialise them.
 * The FIRST call on each device looks up the array's device address (and loads the kernels); make that first call
 * during start-up, before any stream capture begins.  No cudaMalloc / cudaFree / cudaHostAlloc or other allocating
 * routine is called, on the first call or later.  Same decomposition and summation order as gdy_energy on one GPU.
 * Errors are reported through gdy_last_error(NULL). */
#define GDY_NOALLOC_MAX_N 1048576u
int gdy_energy_noalloc(const double *d_r4, const double *d_v4, uint32_t n, double *d_energy_out, void *stream);

/* Distance of every particle to its k-th nearest neighbour (1 <= k <= 8) at the current positions:
 * the O(N^2) part of NbodyUtils::get_center_of_density (reference: src/include/utils/NbodyUtils.cpp:
 * 163-220, which sorts all N distances per particle on the host and takes d[J], the zero self entry sorted
 * first).  dist_out = n doubles with the bits of the reference's d[k] (r^2 in the reference's operation order, no
 * contraction, IEEE sqrt).  n <= k is GDY_ERR_INVALID. */
int gdy_neighbour_distance(gdy_ctx *ctx, int k, double *dist_out);
/* The same kernel bracketed by CUDA events, nothing downloaded (bench: FP64 roofline of k_knn_distance). */
int gdy_neighbour_timed(gdy_ctx *ctx, int k, float *ms_out);

/* ---- fused drivers (same arithmetic as the hooks, fewer launches / host round trips) -------------- */
typedef struct gdy_run_stats
{
    uint64_t iterations;   /* block steps done in this call          */
    uint64_t nsteps;       /* sum of nact                            */
    uint64_t interactions; /* sum of nact * n, 64-bit                */
    double itime;          /* ITIME of the last completed block step */
    double atime;          /* next block time                        */
    double force_ms;       /* device time in the force kernels       */
    double wall_ms;        /* host wall time of the call             */
    uint32_t last_nact;    /* nact of the last completed block step  */
    uint32_t reserved;
} gdy_run_stats;

/* Runs the reference loop body (find -> save_old -> predict -> update -> correct -> next time)
 * until ITIME >= t_stop or max_steps block steps (0 = unlimited).  atime = the next block time
 * (from gdy_init_dt or the previous call).  trace_nact/trace_itime/trace_us (each may be NULL) receive
 * nact, ITIME and the host wall time in microseconds of the first trace_cap block steps. */
int gdy_run(gdy_ctx *ctx, double atime, double itime0, double t_stop, uint64_t max_steps,
            gdy_run_stats *stats, uint32_t *trace_nact, double *trace_itime, float *trace_us,
            uint64_t trace_cap);

/* ---- bench / test entry points ---------------------------------------------------------------------- */
/* Forces of an explicit active list on the current PREDICTED positions (call predicted_pos_vel or
 * init first).  move = nact indices; f6_out (may be NULL) = nact x 6.  *ms_out (may be NULL) = device
 * time of the force kernels (CUDA events on the launching stream). */
int gdy_force_sweep(gdy_ctx *ctx, uint32_t nact, const uint32_t *move, double *f6_out,
                    float *ms_out);
/* Same, nact = n, identity list, no host traffic: `reps` back-to-back evaluations, mean ms. */
int gdy_force_sweep_full(gdy_ctx *ctx, int reps, float *ms_out);

/* gdy_predicted_pos_vel bracketed by CUDA events (bench: GB/s of the memory-bound predictor kernel;
 * replaces k_prediction, reference: src/include/gpu/Hermite4GPU_kernels.cu:127-158). */
int gdy_predict_timed(gdy_ctx *ctx, double itime, float *ms_out);

/* Force-kernel variant override for tuning (-1 = heuristic). */
int gdy_set_tuning(gdy_ctx *ctx, int variant, int jsplit);

/* Small active sets (nact <= nact_max) are stepped by ONE persistent cooperative kernel per run of consecutive small
 * blocks inside gdy_run, without the host (replaces, for those steps, the per-step launch sequence of the reference:
 * src/include/gpu/Hermite4GPU_integration.cpp:81-124).  Every CTA owns a slice of the particles, keeps its raw state
 * resident in shared memory and talks to the other CTAs through flagged slots only: no grid barrier, no fence
 * (csrc/gdy_persist2.cuh).  Every WAIT in it is bounded by a watchdog (environment GDY_SPIN_TIMEOUT_MS, default 2000,
 * read at gdy_create): the timer starts when a wait turns slow and is cleared when the wait succeeds, so a launch may
 * run for any length of time, but a dead or missing peer makes gdy_run return GDY_ERR_STATE instead of hanging.  The
 * same bound applies to the first exchange after gdy_run is entered: a rank that enters gdy_run more than the timeout
 * after its peers (e.g. because it alone wrote a snapshot) is taken for dead; drivers with such skew raise the value.
 * nact_max: 0 = off, < 0 = default, at most 2048.  Sharded contexts need gdy_peer_attach first. */
int gdy_set_persist(gdy_ctx *ctx, int nact_max);
int gdy_persist_stats(const gdy_ctx *ctx, uint64_t *launches_out, uint64_t *steps_out);
/* Can the persistent kernel serve this context?  Returns 0 = switched off (gdy_set_persist(ctx, 0)), 1 = yes,
 * 2 = NO: a CTA's slice of the predicted set (7 doubles per particle of the j-shard / number of SMs) plus its slice of
 * t+dt (1 double per particle of ALL n / number of SMs) plus ~70 KB of fixed tables exceed the 227 KB of shared memory
 * of an SM — above roughly 373 000 particles on one 148-SM GPU (N = 512k or 1M on one GPU; N = 1M fits from four GPUs
 * on).  Such a context is NOT an error: gdy_run steps its small blocks through the per-hook kernels (about 50 us
 * instead of 15 us per small block step), gdy_persist_stats stays at 0, and this call is how a driver finds out.
 * smem_need_out / smem_max_out (may be NULL): bytes needed per CTA and the device's limit.  Negative = gdy_status. */
int gdy_persist_state(gdy_ctx *ctx, uint32_t *smem_need_out, uint32_t *smem_max_out);
/* ns per phase of that kernel seen by CTA 0 (slots 0-8: load, publish, predict+scan, wait for the active particles,
 * pair sums, owners' correction, wait for the corrected, lists, exit; 12/13: summed exchange wait of the owner warps
 * and their count); all zero unless GDY_PERSIST_PROF=1 was in the environment at gdy_create (measurement aid). */
int gdy_persist_profile(gdy_ctx *ctx, uint64_t *ns_out /* [144] */);

/* Peer memory for the sharded paths.  With world > 1 the per-shard force sums are exchanged without NCCL for every
 * block of up to 16384 active particles: every rank stores its sums as flagged 16-byte slots straight into every
 * peer's buffer over NVLink and adds the world's sums in rank order (bit-identical on all ranks) — inside the
 * persistent kernel for small blocks, as k_xsend + a receive fused into the corrector for mid blocks (replaces the
 * host-side sum over GPUs of the reference, src/include/gpu/Hermite4GPU.cu:512-530; bigger blocks keep
 * ncclAllReduce).
 *   gdy_peer_export: handle of THIS context's exchange buffer (GDY_PEER_HANDLE_BYTES bytes);
 *   gdy_peer_attach: all `world` handles in rank order, gathered by the caller (MPI, torch.distributed,
 *                    a vector in the one-process-many-threads host program).  Same process: peer access (same
 *                    device: the same address space — two logical ranks on one GPU, used by the tests);
 *                    other processes: CUDA IPC.  Switches the persistent path on for this shard. */
#define GDY_PEER_HANDLE_BYTES 128
int gdy_peer_export(gdy_ctx *ctx, uint8_t *handle_out);
int gdy_peer_attach(gdy_ctx *ctx, const uint8_t *handles);

/* Measured denominators for the roofline (measurement infrastructure, not the hot path):
 * dependent-free DFMA loop on every SM -> TFLOP/s (2 flop per DFMA), and a device copy -> GB/s. */
int gdy_measure_fp64_peak(gdy_ctx *ctx, double *tflops_out, double *sm_mhz_out);
int gdy_measure_copy_bw(gdy_ctx *ctx, double *gbs_out);

/* The launch plan gdy_update_acc_jrk / gdy_init_acc_jrk would use for `nact` active particles of an n-particle
 * shard (rank, world) on a GPU with num_sms SMs and resident_ctas force CTAs per SM (<= 0: 4), computed on the
 * host without touching a device (tests: every j-unit of the shard is covered exactly once, the partial buffer
 * holds all j-splits).  out[8] = {variant, kind 0 warp / 1 lane, i-tiles, j-splits, j-units per split, first
 * j-unit, end j-unit, ld}; j-units = blocks of 64 particles (warp kernels) or particles (lane kernels). */
int gdy_plan_dry(uint32_t n, uint32_t nact, int rank, int world, int num_sms, int resident_ctas, uint32_t *out);

/* Number of kernels this context has launched (bench.py's gpu_launches). */
uint64_t gdy_launch_count(const gdy_ctx *ctx);
/* How many force exchanges between the shards went through peer memory (k_xsend + the corrector's flagged-slot
 * receive, blocks of <= 16384 active particles once gdy_peer_attach has run) and how many through ncclAllReduce. */
int gdy_exchange_stats(const gdy_ctx *ctx, uint64_t *peer_out, uint64_t *nccl_out);
int gdy_sync(gdy_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* GRAVIDY_B200_H */
