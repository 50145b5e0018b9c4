// gdy_step.cuh — the rest of the block-timestep step on the device: predictor, save-old, partial-force
// reduction, corrector + timestep levels, scheduler (next block time, ordered active-list compaction),
// energy, AoS<->SoA packing, and the two measurement micro-kernels.
//
// Everything that decides the SCHEDULE (predictor, corrector, timestep criterion, level snapping, the
// active-set test) is written with explicit round-to-nearest intrinsics (__dadd_rn, __dmul_rn,
// __ddiv_rn, __dsqrt_rn): nvcc never contracts those into FMAs, so given identical forces the results
// are bit-identical to the reference's CPU back end, which is compiled without FMA (reference:
// src/Makefile:40-54) — the GPU back end of the reference uses different constants and is NOT followed
// (reference: src/include/gpu/Hermite4GPU_kernels.cu:141-142, src/include/gpu/Hermite4GPU.cu:293-298).
#pragma once
#include "gdy_device.cuh"

namespace gdy {

constexpr double D_TIME_MIN = 1.1920928955078125e-07;  // 2^-23, reference: src/include/common.hpp:120
constexpr double D_TIME_MAX = 0.125;                   // 2^-3,  reference: src/include/common.hpp:126
constexpr double DBL_EPS2 = 2.0 * 2.220446049250313e-16;  // 2*epsilon, reference: Hermite4.cpp:76-77

__device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double dvd(double a, double b) { return __ddiv_rn(a, b); }

// Taylor predictor of one particle.  Expression order of Hermite4CPU::predicted_pos_vel
// (reference: src/include/cpu/Hermite4CPU.cpp:346-359): dt2 = dt*dt, dt3 = dt2*dt, dt2/2.0, dt3/6.0,
// ((dt3c*j + dt2c*a) + dt*v) + r and (dt2c*j + dt*a) + v.
struct Pred6 { double x, y, z, vx, vy, vz; };
__device__ __forceinline__ Pred6 predict_vals(double itime, double t, double x, double y, double z, double vx,
                                              double vy, double vz, double ax, double ay, double az,
                                              double jx, double jy, double jz)
{
    const double dt = sub(itime, t);
    const double dt2 = mul(dt, dt);
    const double dt3 = mul(dt2, dt);
    const double dt2c = dvd(dt2, 2.0);
    const double dt3c = dvd(dt3, 6.0);
    Pred6 p;
    p.x = add(add(add(mul(dt3c, jx), mul(dt2c, ax)), mul(dt, vx)), x);
    p.y = add(add(add(mul(dt3c, jy), mul(dt2c, ay)), mul(dt, vy)), y);
    p.z = add(add(add(mul(dt3c, jz), mul(dt2c, az)), mul(dt, vz)), z);
    p.vx = add(add(mul(dt2c, jx), mul(dt, ax)), vx);
    p.vy = add(add(mul(dt2c, jy), mul(dt, ay)), vy);
    p.vz = add(add(mul(dt2c, jz), mul(dt, az)), vz);
    return p;
}
__device__ __forceinline__ Pred6 predict_one(const State &s, uint32_t i, double itime)
{
    return predict_vals(itime, s.t[i], s.x[i], s.y[i], s.z[i], s.vx[i], s.vy[i], s.vz[i], s.f[0][i], s.f[1][i],
                        s.f[2][i], s.f[3][i], s.f[4][i], s.f[5][i]);
}

__device__ __forceinline__ void store_pred(const State &s, uint32_t i, const Pred6 &p)
{
    double *q = s.pb + pbidx(i, 0);
    q[0] = p.x; q[JB] = p.y; q[2 * JB] = p.z;
    q[3 * JB] = p.vx; q[4 * JB] = p.vy; q[5 * JB] = p.vz;
}

// predicted_pos_vel(ITIME) for particles [begin, end), written to the blocked predicted array.
// Replaces k_prediction (reference: src/include/gpu/Hermite4GPU_kernels.cu:127-158).
// nextra > 0: threads past the range predict the active particles move[0..nextra) as well — with the
// j-particles sharded over GPUs a rank predicts only its own j-slice, but the force kernel reads the
// predicted i-particles and the corrector needs p[i] of EVERY active particle.  Same arithmetic, so a
// particle predicted twice gets the same bits twice.
__global__ void k_predict(State s, uint32_t begin, uint32_t end, double itime,
                          const uint32_t *__restrict__ move, uint32_t nextra)
{
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t i;
    if (g < end - begin) i = begin + g;
    else if (g - (end - begin) < nextra) i = move[g - (end - begin)];
    else return;
    store_pred(s, i, predict_one(s, i, itime));
}

// Predicted set of the particles that were NOT corrected at `itime` (t != itime): used after the
// persistent small-step kernel, which predicts in registers and stores only the active particles.
__global__ void k_predict_rest(State s, uint32_t n, double itime)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || s.t[i] == itime) return;
    store_pred(s, i, predict_one(s, i, itime));
}

// init: predicted = current (reference: Hermite4::init_data, src/include/Hermite4.cpp:198-210)
__global__ void k_copy_to_pred(State s, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    store_pred(s, i, Pred6{s.x[i], s.y[i], s.z[i], s.vx[i], s.vy[i], s.vz[i]});
}

// mass row (and, for the padding particles, the far-away position) of the blocked predicted array
__global__ void k_init_pb(State s, uint32_t n, uint32_t npad)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npad) return;
    double *q = s.pb + pbidx(i, 0);
    if (i < n)
    {
        // The reference keeps every mass as a float: the reader narrows it (NbodySystem.cpp:174) and so does Predictor.m
        // (common.hpp:216), which is what the force loop reads.  A caller of the C ABI that hands over a mass that is
        // not float-representable gets the same narrowing here, once, for every array that holds the mass.
        const double mf = (double)(float)s.m[i];
        s.m[i] = mf;
        q[6 * JB] = mf;
    }
    else
    {
        q[0] = PAD_POS; q[JB] = PAD_POS; q[2 * JB] = PAD_POS;
        q[3 * JB] = 0.0; q[4 * JB] = 0.0; q[5 * JB] = 0.0; q[6 * JB] = 0.0;
    }
}

// save_old_acc_jrk(nact) (reference: src/include/Hermite4.cpp:145-153)
__global__ void k_save_old(State s, uint32_t nact)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nact) return;
    const uint32_t i = s.move[k];
#pragma unroll
    for (int c = 0; c < 6; ++c) s.old[c][i] = s.f[c][i];
}

// h_f[move[k]] = summed force (the tail of update_acc_jrk).  fbuf = [6][ld], k-indexed.
__global__ void k_scatter_forces(State s, const uint32_t *__restrict__ move, uint32_t nact,
                                 const double *__restrict__ fbuf, uint32_t ld)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nact) return;
    const uint32_t i = move ? move[k] : k;
#pragma unroll
    for (int c = 0; c < 6; ++c) s.f[c][i] = fbuf[(size_t)c * ld + k];
}

// ---- corrector + timestep ---------------------------------------------------------------------
__device__ __forceinline__ double magnitude(double x, double y, double z)
{
    // reference: NbodyUtils::get_magnitude, src/include/utils/NbodyUtils.cpp:331-334
    return __dsqrt_rn(add(add(mul(x, x), mul(y, y)), mul(z, z)));
}

// reference: NbodyUtils::normalize_dt, src/include/utils/NbodyUtils.cpp:371-424
__device__ __forceinline__ double normalize_dt(double new_dt, double old_dt, double t)
{
    const double o8 = dvd(old_dt, 8.0), o4 = dvd(old_dt, 4.0), o2 = dvd(old_dt, 2.0);
    if (new_dt <= o8) new_dt = D_TIME_MIN;
    else if (o8 < new_dt && new_dt <= o4) new_dt = o8;
    else if (o4 < new_dt && new_dt <= o2) new_dt = o4;
    else if (o2 < new_dt && new_dt <= old_dt) new_dt = o2;
    else if (old_dt < new_dt && new_dt <= mul(old_dt, 2.0)) new_dt = old_dt;
    else if (mul(2.0, old_dt) < new_dt)
    {
        const double val = dvd(t, mul(2.0, old_dt));
        new_dt = (ceil(val) == val) ? mul(2.0, old_dt) : old_dt;
    }
    else new_dt = old_dt;   // NaN falls through here, as in the reference

    if (new_dt < D_TIME_MIN) new_dt = D_TIME_MIN;
    else if (new_dt > D_TIME_MAX) new_dt = D_TIME_MAX;
    return new_dt;
}

// correction_pos_vel(ITIME, nact) with get_timestep_normal + normalize_dt.  Expression order of the
// CPU back end (reference: src/include/cpu/Hermite4CPU.cpp:370-411, NbodyUtils.cpp:337-365).
// The reference runs this on the host even in its GPU build (reference: Hermite4GPU.cu:274-328).
// fbuf != NULL (fused driver): the new forces are taken from the k-indexed buffer [6][ld] the force path
// produced and stored to h_f here, and the forces they replace are saved to h_old on the way — the
// scatter at the end of update_acc_jrk and save_old_acc_jrk (reference: src/include/Hermite4.cpp:145-153)
// fused into the corrector; same values, two launches fewer.
// The arithmetic of one particle: old forces (oa, oj), new forces (fa, fj), old step dt1, predicted p.
// Writes a2, a3, r, v, t, dt of particle i and returns the corrected r, v and the new step.
struct CorrIn { double dt1; double oa[3], oj[3], fa[3], fj[3]; Pred6 p; };
struct CorrOut { double x, y, z, vx, vy, vz, ndt; };
__device__ __forceinline__ CorrOut correct_core(const State &s, uint32_t i, double itime, double eta, const CorrIn &q)
{
    const double dt1 = q.dt1;
    const double dt2 = mul(dt1, dt1);
    const double dt3 = mul(dt2, dt1);
    const double dt4 = mul(dt2, dt2);
    const double dt5 = mul(dt4, dt1);
    // Every step the scheduler hands out is a power of two (init_dt: pow(2, k); normalize_dt: old/8, old/4,
    // old/2, old, 2*old, 2^-23, 2^-3).  Dividing BY a power of two, or a power of two by a constant, is an
    // exact scaling of a correctly rounded constant, so the six divisions below are done as multiplications
    // with bit-identical results (no subnormals: dt >= 2^-23); anything else takes the true divisions.
    const bool pow2 = (__double_as_longlong(dt1) & 0x000fffffffffffffLL) == 0 && dt1 >= D_TIME_MIN && dt1 <= 1.0;
    double c4_24, c5_120, c3_6, inv_dt2 = 0.0, inv_dt3 = 0.0;
    if (pow2)
    {
        c4_24 = mul(dt4, dvd(1.0, 24.0)); c5_120 = mul(dt5, dvd(1.0, 120.0)); c3_6 = mul(dt3, dvd(1.0, 6.0));
        // 1/dt2, 1/dt3 by exponent arithmetic: dt^k = 2^e  ->  bits(2^-e) = (2046 << 52) - bits(2^e)
        inv_dt2 = __longlong_as_double((2046LL << 52) - __double_as_longlong(dt2));
        inv_dt3 = __longlong_as_double((2046LL << 52) - __double_as_longlong(dt3));
    }
    else { c4_24 = dvd(dt4, 24.0); c5_120 = dvd(dt5, 120.0); c3_6 = dvd(dt3, 6.0); }

    double a2[3], a3[3];
#pragma unroll
    for (int c = 0; c < 3; ++c)
    {
        const double d = sub(q.oa[c], q.fa[c]);
        // (-6*(old.a - f.a) - dt1*(4*old.a1 + 2*f.a1)) / dt2
        const double n2 = sub(mul(-6.0, d), mul(dt1, add(mul(4.0, q.oj[c]), mul(2.0, q.fj[c]))));
        // (12*(old.a - f.a) + 6*dt1*(old.a1 + f.a1)) / dt3
        const double n3 = add(mul(12.0, d), mul(mul(6.0, dt1), add(q.oj[c], q.fj[c])));
        a2[c] = pow2 ? mul(n2, inv_dt2) : dvd(n2, dt2);
        a3[c] = pow2 ? mul(n3, inv_dt3) : dvd(n3, dt3);
        s.a2[c][i] = a2[c];
        s.a3[c][i] = a3[c];
    }
    // r = p.r + (dt4/24)*a2 + (dt5/120)*a3 ;  v = p.v + (dt3/6)*a2 + (dt4/24)*a3
    CorrOut o;
    o.x = add(add(q.p.x, mul(c4_24, a2[0])), mul(c5_120, a3[0]));
    o.y = add(add(q.p.y, mul(c4_24, a2[1])), mul(c5_120, a3[1]));
    o.z = add(add(q.p.z, mul(c4_24, a2[2])), mul(c5_120, a3[2]));
    o.vx = add(add(q.p.vx, mul(c3_6, a2[0])), mul(c4_24, a3[0]));
    o.vy = add(add(q.p.vy, mul(c3_6, a2[1])), mul(c4_24, a3[1]));
    o.vz = add(add(q.p.vz, mul(c3_6, a2[2])), mul(c4_24, a3[2]));
    s.x[i] = o.x; s.y[i] = o.y; s.z[i] = o.z;
    s.vx[i] = o.vx; s.vy[i] = o.vy; s.vz[i] = o.vz;
    s.t[i] = itime;

    // get_timestep_normal(i, eta): h_dt[i] is still the old step here
    const double ax1_2 = add(a2[0], mul(dt1, a3[0]));
    const double ay1_2 = add(a2[1], mul(dt1, a3[1]));
    const double az1_2 = add(a2[2], mul(dt1, a3[2]));
    const double abs_a1 = magnitude(q.fa[0], q.fa[1], q.fa[2]);
    const double abs_j1 = magnitude(q.fj[0], q.fj[1], q.fj[2]);
    const double abs_j12 = mul(abs_j1, abs_j1);
    const double abs_a1_3 = magnitude(a3[0], a3[1], a3[2]);
    const double abs_a1_2 = magnitude(ax1_2, ay1_2, az1_2);
    const double abs_a1_22 = mul(abs_a1_2, abs_a1_2);
    const double normal_dt = __dsqrt_rn(
        mul(eta, dvd(add(mul(abs_a1, abs_a1_2), abs_j12), add(mul(abs_j1, abs_a1_3), abs_a1_22))));
    o.ndt = normalize_dt(normal_dt, dt1, itime);
    s.dt[i] = o.ndt;
    return o;
}

// ---- one-shot exchange of the ranks' partial forces through peer memory (blocks of <= XMID_LD active particles) ----
// Replaces ncclAllReduce (and the reference's host-side sum over GPUs, src/include/gpu/Hermite4GPU.cu:512-530) for
// the blocks that are too big for the persistent kernel and too small for a bandwidth-optimal collective: k_xsend
// stores this rank's sums as flagged 16-byte slots {lo, tag, hi, tag} into buffer [parity][my rank] of EVERY rank
// over NVLink; the corrector of every rank polls its own buffer and adds the world's sums in rank order, so all
// ranks correct with the same bits.  No barrier, no second collective kernel.  Two parities: a rank reaches the
// send of exchange e+1 only behind its corrector of exchange e, which needed every rank's send of e, which they
// issue (stream order) behind their corrector of e-1 — the last reader of the parity being rewritten.
constexpr int XMID_LD = 16384;            // active particles per exchange at most
constexpr int XMID_MAXW = 8;
struct XRecv
{
    const uint4 *base;                    // my exchange buffer: [2][XMID_MAXW][6][XMID_LD] slots (NULL: fbuf holds the sums)
    int world;
    uint32_t tag, par;
    uint32_t *abort_host;                 // pinned: set when a poll gives up
    unsigned long long timeout_ns;
};
// XSend and xll_store_sys: gdy_force.cuh (k_reduce_partials sends as well)
static_assert(XMID_MAXW == 8, "XSend holds eight peers");

__global__ void __launch_bounds__(256)
k_xsend(const double *__restrict__ fbuf, uint32_t ld, uint32_t nact, XSend x)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nact) return;
#pragma unroll
    for (int c = 0; c < 6; ++c)
    {
        const double v = fbuf[(size_t)c * ld + k];
        const size_t at = (((size_t)x.par * XMID_MAXW + x.rank) * 6 + c) * XMID_LD + k;
        for (int r = 0; r < x.world; ++r) xll_store_sys(x.peer[r] + at, v, x.tag);
    }
}

__device__ __forceinline__ double xrecv_sum(const XRecv &x, int c, uint32_t k)
{
    double sum = 0.0;
    uint4 v[XMID_MAXW];
    unsigned long long t0 = 0;
    uint32_t spins = 0;
    bool ok;
    do {
        ok = true;
#pragma unroll
        for (int r = 0; r < XMID_MAXW; ++r)
            if (r < x.world)
            {
                const uint4 *p = x.base + (((size_t)x.par * XMID_MAXW + r) * 6 + c) * XMID_LD + k;
                asm volatile("ld.relaxed.sys.global.v4.b32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(v[r].x), "=r"(v[r].y), "=r"(v[r].z), "=r"(v[r].w) : "l"(p) : "memory");
                ok = ok && v[r].y == x.tag && v[r].w == x.tag;
            }
        if (!ok && (++spins & 255u) == 0)
        {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            else if (t - t0 > x.timeout_ns || *(volatile uint32_t *)x.abort_host)
            {
                *(volatile uint32_t *)x.abort_host = 1u;   // the host fails the step (no reference behaviour to keep: it exits)
                __threadfence_system();
                return 0.0;
            }
        }
    } while (!ok);
#pragma unroll
    for (int r = 0; r < XMID_MAXW; ++r)
        if (r < x.world)
            sum += __longlong_as_double((long long)(((unsigned long long)v[r].z << 32) | (unsigned long long)v[r].x));
    return sum;
}

// the hook-by-hook path keeps update_acc_jrk and correction_pos_vel apart: the world's sums -> out[6][ld]
__global__ void __launch_bounds__(256)
k_xrecv(double *__restrict__ out, uint32_t ld, uint32_t nact, XRecv xr)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nact) return;
#pragma unroll
    for (int c = 0; c < 6; ++c) out[(size_t)c * ld + k] = xrecv_sum(xr, c, k);
}

__global__ void k_correct(State s, uint32_t nact, double itime, double eta,
                          const double *__restrict__ fbuf, uint32_t ld, XRecv xr)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nact) return;
    const uint32_t i = s.move[k];
    CorrIn q;
    q.dt1 = s.dt[i];
#pragma unroll
    for (int c = 0; c < 3; ++c)
    {
        if (xr.base)
        {
            q.oa[c] = s.f[c][i]; q.oj[c] = s.f[3 + c][i];
            s.old[c][i] = q.oa[c]; s.old[3 + c][i] = q.oj[c];
            q.fa[c] = xrecv_sum(xr, c, k);
            q.fj[c] = xrecv_sum(xr, 3 + c, k);
            s.f[c][i] = q.fa[c];
            s.f[3 + c][i] = q.fj[c];
        }
        else if (fbuf)
        {
            q.oa[c] = s.f[c][i]; q.oj[c] = s.f[3 + c][i];
            s.old[c][i] = q.oa[c]; s.old[3 + c][i] = q.oj[c];
            q.fa[c] = fbuf[(size_t)c * ld + k];
            q.fj[c] = fbuf[(size_t)(3 + c) * ld + k];
            s.f[c][i] = q.fa[c];
            s.f[3 + c][i] = q.fj[c];
        }
        else
        {
            q.oa[c] = s.old[c][i]; q.oj[c] = s.old[3 + c][i];
            q.fa[c] = s.f[c][i];
            q.fj[c] = s.f[3 + c][i];
        }
    }
    const double *pp = s.pb + pbidx(i, 0);
    q.p = Pred6{pp[0], pp[JB], pp[2 * JB], pp[3 * JB], pp[4 * JB], pp[5 * JB]};
    correct_core(s, i, itime, eta, q);
}

// ---- scheduler -----------------------------------------------------------------------------------
// next_integration_time: min over all particles of t+dt (reference: src/include/Hermite4.cpp:90-102),
// in two phases without atomics: k_min_time leaves one minimum per block, the consumer (k_sched_count /
// k_sched_write / k_min_final) folds the <= MIN_BLOCKS block minima itself.  fmin is exact and
// order-independent, so every block (and every GPU) gets the same bit pattern.
constexpr int MIN_BLOCKS = 256;

__device__ __forceinline__ double block_min_256(double m, double *sm /* [8] */)
{
#pragma unroll
    for (int off = 16; off; off >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, off));
    __syncthreads();   // protect sm from a previous use
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    m = sm[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) m = fmin(m, sm[w]);
    return m;
}

__global__ void __launch_bounds__(256)
k_min_time(State s, uint32_t n, double *__restrict__ block_min)
{
    __shared__ double sm[8];
    double m = __longlong_as_double(0x7ff0000000000000LL);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        m = fmin(m, add(s.t[i], s.dt[i]));
    m = block_min_256(m, sm);
    if (threadIdx.x == 0) block_min[blockIdx.x] = m;
}

// fold the block minima (called by all 256 threads of a block)
__device__ __forceinline__ double fold_block_min(const double *__restrict__ block_min, uint32_t nmin,
                                                 double *sm)
{
    double m = __longlong_as_double(0x7ff0000000000000LL);
    if (threadIdx.x < nmin) m = block_min[threadIdx.x];
    return block_min_256(m, sm);
}

__global__ void __launch_bounds__(256)
k_min_final(const double *__restrict__ block_min, uint32_t nmin, double *atime_dev, volatile double *atime_host,
            volatile uint32_t *seq_host, uint32_t seq)
{
    __shared__ double sm[8];
    const double m = fold_block_min(block_min, nmin, sm);
    if (threadIdx.x == 0)
    {
        *atime_dev = m;
        if (atime_host)
        {
            *atime_host = m;
            __threadfence_system();
            *seq_host = seq;   // written last: the host polls on it
        }
    }
}

// find_particles_to_move(ITIME): ordered compaction, two passes over chunks of 1024 particles.
// Active test of the reference: fabs(ITIME - (t+dt)) < 2*epsilon (src/include/Hermite4.cpp:75-77).
// block_min != NULL: ITIME is the minimum left by k_min_time (the next block time), so the scheduler
// of step k+1 chains behind the corrector of step k without a host round trip.
constexpr int SCHED_ITEMS = 4;
constexpr int SCHED_BLOCK = 256;
constexpr int SCHED_CHUNK = SCHED_ITEMS * SCHED_BLOCK;

__device__ __forceinline__ bool is_active(const State &s, uint32_t i, uint32_t n, double itime)
{
    if (i >= n) return false;
    const double tmp_time = add(s.t[i], s.dt[i]);
    return fabs(sub(itime, tmp_time)) < DBL_EPS2;
}

__global__ void __launch_bounds__(SCHED_BLOCK)
k_sched_count(State s, uint32_t n, double itime, const double *__restrict__ block_min, uint32_t nmin,
              uint32_t *__restrict__ block_cnt)
{
    __shared__ double sm[8];
    __shared__ int wsum[SCHED_BLOCK / 32];
    if (block_min) itime = fold_block_min(block_min, nmin, sm);
    const uint32_t base = blockIdx.x * SCHED_CHUNK + threadIdx.x * SCHED_ITEMS;
    int v = 0;
#pragma unroll
    for (int e = 0; e < SCHED_ITEMS; ++e) v += is_active(s, base + e, n, itime) ? 1 : 0;
#pragma unroll
    for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        int b = 0;
#pragma unroll
        for (int w = 0; w < SCHED_BLOCK / 32; ++w) b += wsum[w];
        block_cnt[blockIdx.x] = (uint32_t)b;
    }
}

// out2_host (pinned, may be NULL) receives {ITIME bits, nact}: the one readback per block step.
struct SchedOut { double itime; uint32_t nact; uint32_t seq; };

__global__ void __launch_bounds__(SCHED_BLOCK)
k_sched_write(State s, uint32_t n, double itime, const double *__restrict__ block_min, uint32_t nmin,
              const uint32_t *__restrict__ block_cnt, uint32_t nblocks, SchedOut *out_dev,
              volatile SchedOut *out_host, uint32_t seq)
{
    __shared__ double sm[8];
    __shared__ uint32_t sh[SCHED_BLOCK / 32];
    __shared__ uint32_t wb[SCHED_BLOCK / 32], wt[SCHED_BLOCK / 32];
    if (block_min) itime = fold_block_min(block_min, nmin, sm);
    // (1) offset of this block = sum of the counts of the blocks before it (and the grand total)
    uint32_t before = 0, total = 0;
    for (uint32_t b = threadIdx.x; b < nblocks; b += SCHED_BLOCK)
    {
        const uint32_t c = block_cnt[b];
        total += c;
        if (b < blockIdx.x) before += c;
    }
#pragma unroll
    for (int off = 16; off; off >>= 1)
    {
        before += __shfl_xor_sync(0xffffffffu, before, off);
        total += __shfl_xor_sync(0xffffffffu, total, off);
    }
    if ((threadIdx.x & 31) == 0) { wb[threadIdx.x >> 5] = before; wt[threadIdx.x >> 5] = total; }
    // (2) flags of my items, exclusive scan inside the block, ordered write
    const uint32_t base = blockIdx.x * SCHED_CHUNK + threadIdx.x * SCHED_ITEMS;
    bool fl[SCHED_ITEMS];
    uint32_t c = 0;
#pragma unroll
    for (int e = 0; e < SCHED_ITEMS; ++e) { fl[e] = is_active(s, base + e, n, itime); c += fl[e] ? 1u : 0u; }
    uint32_t incl = c;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1)
    {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += v;
    }
    if (lane == 31) sh[w] = incl;
    __syncthreads();
    uint32_t s_off = 0, s_tot = 0;
#pragma unroll
    for (int ww = 0; ww < SCHED_BLOCK / 32; ++ww) { s_off += wb[ww]; s_tot += wt[ww]; }
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        out_dev->itime = itime; out_dev->nact = s_tot; out_dev->seq = seq;
        if (out_host)
        {
            out_host->itime = itime; out_host->nact = s_tot;
            __threadfence_system();
            out_host->seq = seq;   // written last: the host may poll on it
        }
    }
    uint32_t woff = 0;
    for (int ww = 0; ww < w; ++ww) woff += sh[ww];
    uint32_t pos = s_off + woff + incl - c;
#pragma unroll
    for (int e = 0; e < SCHED_ITEMS; ++e)
        if (fl[e]) s.move[pos++] = base + e;
}

// ---- energy ------------------------------------------------------------------------------------------
// Kinetic 0.5 m v^2 and UNSOFTENED potential -sum_{i<j} m_i m_j / r_ij (reference:
// src/include/utils/NbodyUtils.cpp:427-529; k_energy, src/include/gpu/Hermite4GPU_kernels.cu:250-288).
// The reference's triangular j>i loop is load-imbalanced; here every i sums over all j != i and the result is
// halved.  A CTA of 256 threads holds EN_CHUNK = 512 i-particles (two per thread, in registers) and streams its
// j-range through double-buffered shared-memory tiles of 256 particles (rows x y z m, read as broadcast LDS.128 = two
// j per load): 12 FP64-pipe instructions + 1 compare per ordered pair.  Grid = (i-chunks of this rank, j-splits).
// Every (chunk, split) sum is produced by one CTA in a fixed order and the chunk sums are added in chunk order, so
// the energy has the same bits whatever the number of GPUs the chunks are spread over (reference: i-slices per GPU,
// src/include/gpu/Hermite4GPU.cu:540-592).
constexpr int EN_T = 256;
constexpr int EN_CHUNK = 2 * EN_T;
constexpr int EN_TILE = 256;

__device__ __forceinline__ double block_sum_256(double v, double *sm /* [8] */)
{
#pragma unroll
    for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = sm[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) t += sm[w];
    return t;
}

__global__ void __launch_bounds__(EN_T)
k_energy(State s, uint32_t n, uint32_t chunk0, uint32_t tiles_per_split, uint32_t ntiles,
         double *__restrict__ cta_ek /* [nch] */, double *__restrict__ cta_ep /* [nsplit][nch_total] */, uint32_t nch_total)
{
    __shared__ __align__(16) double tile[2][4][EN_TILE];
    __shared__ double red[8];
    const int tid = threadIdx.x;
    const uint32_t chunk = chunk0 + blockIdx.x;
    const uint32_t i0 = chunk * EN_CHUNK + tid, i1 = i0 + EN_T;
    const uint32_t c0 = i0 < n ? i0 : n - 1, c1 = i1 < n ? i1 : n - 1;
    const double x0 = s.x[c0], y0 = s.y[c0], z0 = s.z[c0];
    const double x1 = s.x[c1], y1 = s.y[c1], z1 = s.z[c1];
    const uint32_t t0 = blockIdx.y * tiles_per_split, t1 = min(t0 + tiles_per_split, ntiles);
    double p0 = 0.0, p1 = 0.0;
    auto stage = [&](uint32_t t, int b) {
        const uint32_t j = t * EN_TILE + tid;   // < npad: padding particles have zero mass and sit far away
        tile[b][0][tid] = s.x[j]; tile[b][1][tid] = s.y[j]; tile[b][2][tid] = s.z[j]; tile[b][3][tid] = s.m[j];
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
                y = (r2 > 0.0) ? y : 0.0;   /* i == j (and coincident particles: the reference gives inf there) */ \
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
    double ep = 0.0, ek = 0.0;
    if (i0 < n) ep += -0.5 * s.m[i0] * p0;
    if (i1 < n) ep += -0.5 * s.m[i1] * p1;
    ep = block_sum_256(ep, red);
    if (tid == 0) cta_ep[(size_t)blockIdx.y * nch_total + chunk] = ep;
    if (blockIdx.y == 0)
    {
        if (i0 < n) { const double vx = s.vx[i0], vy = s.vy[i0], vz = s.vz[i0]; ek += 0.5 * s.m[i0] * (vx * vx + vy * vy + vz * vz); }
        if (i1 < n) { const double vx = s.vx[i1], vy = s.vy[i1], vz = s.vz[i1]; ek += 0.5 * s.m[i1] * (vx * vx + vy * vy + vz * vz); }
        ek = block_sum_256(ek, red);
        if (tid == 0) cta_ek[chunk] = ek;
    }
}

// chunk sums: the j-splits of a chunk in split order.  out = [2][nch_total]: kinetic, potential (chunks of other ranks: 0)
__global__ void k_energy_fold(const double *__restrict__ cta_ek, const double *__restrict__ cta_ep, uint32_t nsplit,
                              uint32_t nch_total, uint32_t chunk0, uint32_t nch, double *__restrict__ out)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nch_total) return;
    double ek = 0.0, ep = 0.0;
    if (c >= chunk0 && c < chunk0 + nch)
    {
        ek = cta_ek[c];
        for (uint32_t sp = 0; sp < nsplit; ++sp) ep += cta_ep[(size_t)sp * nch_total + c];
    }
    out[c] = ek;
    out[nch_total + c] = ep;
}

// ---- k-th nearest neighbour distance ------------------------------------------------------------------
// The O(N^2) part of the Casertano & Hut density centre (reference: NbodyUtils::get_center_of_density,
// src/include/utils/NbodyUtils.cpp:163-220, sorts all N distances of every particle on the host and takes d[J],
// the self entry 0 sorted first).  One thread per particle keeps the KNN_MAX smallest squared distances (self
// excluded) in registers, j through double-buffered shared-memory tiles read as broadcast LDS.128.  r^2 is computed
// WITHOUT contraction in the reference's order, (dx*dx + dy*dy) + dz*dz, so that sqrt of the k-th smallest r^2 has the
// bits of the reference's d[k].  Padding particles sit 1e10 away and enter only when n <= k (refused by the host).
constexpr int KNN_MAX = 8;
constexpr int KNN_T = 128;
__global__ void __launch_bounds__(KNN_T)
k_knn_distance(State s, uint32_t n, uint32_t npad, int k, double *__restrict__ dist)
{
    __shared__ __align__(16) double tile[2][3][KNN_T];
    const int tid = threadIdx.x;
    const uint32_t i = blockIdx.x * KNN_T + tid;
    const uint32_t ic = i < n ? i : n - 1;
    const double xi = s.x[ic], yi = s.y[ic], zi = s.z[ic];
    double best[KNN_MAX];
#pragma unroll
    for (int q = 0; q < KNN_MAX; ++q) best[q] = 1e300;
    auto stage = [&](uint32_t t, int b) {
        tile[b][0][tid] = s.x[t + tid]; tile[b][1][tid] = s.y[t + tid]; tile[b][2][tid] = s.z[t + tid];
    };
    stage(0, 0);
    __syncthreads();
    for (uint32_t t = 0; t < npad; t += KNN_T)
    {
        const int b = (t / KNN_T) & 1;
        if (t + KNN_T < npad) stage(t + KNN_T, b ^ 1);
#pragma unroll 2
        for (int j = 0; j < KNN_T; j += 2)
        {
            const double2 X = *reinterpret_cast<const double2 *>(&tile[b][0][j]);
            const double2 Y = *reinterpret_cast<const double2 *>(&tile[b][1][j]);
            const double2 Z = *reinterpret_cast<const double2 *>(&tile[b][2][j]);
#pragma unroll
            for (int u = 0; u < 2; ++u)
            {
                const double dx = sub(u ? X.y : X.x, xi), dy = sub(u ? Y.y : Y.x, yi), dz = sub(u ? Z.y : Z.x, zi);
                double r2 = add(add(mul(dx, dx), mul(dy, dy)), mul(dz, dz));
                if (t + j + u == ic) r2 = 1e300;            // self
                if (r2 < best[KNN_MAX - 1])
                {
                    // insertion into the sorted list (fully unrolled: stays in registers)
#pragma unroll
                    for (int q = 0; q < KNN_MAX; ++q)
                    {
                        const double lo = fmin(best[q], r2);
                        r2 = fmax(best[q], r2);
                        best[q] = lo;
                    }
                }
            }
        }
        __syncthreads();
    }
    if (i < n)
    {
        double r2 = best[0];
#pragma unroll
        for (int q = 1; q < KNN_MAX; ++q) if (q == k - 1) r2 = best[q];
        dist[i] = __dsqrt_rn(r2);
    }
}

// Deterministic sum of two arrays (fixed tree): out[0] = sum a, out[1] = sum b.  One block.
__global__ void __launch_bounds__(1024) k_sum2(const double *a, const double *b, uint32_t n, double *out)
{
    __shared__ double sa[1024], sb[1024];
    double x = 0.0, y = 0.0;
    for (uint32_t i = threadIdx.x; i < n; i += 1024) { x += a[i]; y += b[i]; }
    sa[threadIdx.x] = x; sb[threadIdx.x] = y;
    __syncthreads();
    for (int off = 512; off; off >>= 1)
    {
        if ((int)threadIdx.x < off) { sa[threadIdx.x] += sa[threadIdx.x + off]; sb[threadIdx.x] += sb[threadIdx.x + off]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { out[0] = sa[0]; out[1] = sb[0]; }
}

// ---- AoS <-> SoA packing ------------------------------------------------------------------------
// aos[i*stride + c] <-> col[c][i]
struct Cols { double *c[8]; };
__global__ void k_unpack(const double *__restrict__ aos, Cols cols, int ncol, int stride, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int c = 0; c < ncol; ++c)
        if (cols.c[c]) cols.c[c][i] = aos[(size_t)i * stride + c];
}
// sharded upload: chunk r of the gather buffer holds `chunk` particles as [chunk][4] positions+mass, then [chunk][4]
// velocities (gdy_upload_particles)
__global__ void k_unpack_chunks(const double *__restrict__ g, Cols rc, Cols vc, uint32_t chunk, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t rk = i / chunk, li = i % chunk;
    const double *base = g + (size_t)rk * chunk * 8;
#pragma unroll
    for (int c = 0; c < 4; ++c) rc.c[c][i] = base[(size_t)li * 4 + c];
#pragma unroll
    for (int c = 0; c < 3; ++c) vc.c[c][i] = base[(size_t)chunk * 4 + (size_t)li * 4 + c];
}
__global__ void k_pack(double *__restrict__ aos, Cols cols, int ncol, int stride, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int c = 0; c < ncol; ++c) aos[(size_t)i * stride + c] = cols.c[c] ? cols.c[c][i] : 0.0;
}
// Predictor h_p as AoS {r[3], v[3], m} from the blocked array (download only)
__global__ void k_pack_pred(double *__restrict__ aos, State s, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *q = s.pb + pbidx(i, 0);
#pragma unroll
    for (int c = 0; c < 7; ++c) aos[(size_t)i * 7 + c] = q[c * JB];
}
__global__ void k_fill(double *p, double v, uint32_t begin, uint32_t end)
{
    const uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) p[i] = v;
}

// ---- measurement micro-kernels (roofline denominators; not on the hot path) -------------------------
// NCH independent DFMA chains per thread: the FP64 pipe's issue rate, nothing else.  The host tries a few
// (chains, residency) shapes and keeps the best, so the denominator is the most the pipe has been seen
// to sustain, not an artefact of one instruction schedule.
template <int NCH>
__global__ void __launch_bounds__(256) k_dfma_peak(double *out, int iters, double a, double b)
{
    double x[NCH];
#pragma unroll
    for (int c = 0; c < NCH; ++c) x[c] = (double)(threadIdx.x + c);
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int u = 0; u < 64 / NCH; ++u)
#pragma unroll
            for (int c = 0; c < NCH; ++c) x[c] = fma(x[c], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < NCH; ++c) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// The same with the accumulator in the addend slot and one per-thread register as both factors,
// x[c] = fma(u[k], u[k], x[c]): two distinct register sources per DFMA.  On B200 this shape retires 63.7 of the
// nominal 64 lanes per clock per SM (the x = fma(x, a, b) shape above stops at 57.8, three distinct register
// sources at 48: tools/fp64_mix.cu, profiles/r01_fp64_mix_v5.md), so it is the one that sets the peak.
template <int NCH>
__global__ void __launch_bounds__(256) k_dfma_peak_acc(double *out, int iters, double a, double b)
{
    double x[NCH], u[8];
#pragma unroll
    for (int c = 0; c < NCH; ++c) x[c] = (double)(threadIdx.x + c);
#pragma unroll
    for (int c = 0; c < 8; ++c) u[c] = a * 1e-3 + b * (double)(threadIdx.x * 8 + c);
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int q = 0; q < 64 / NCH; ++q)
#pragma unroll
            for (int c = 0; c < NCH; ++c) x[c] = fma(u[(c + q) % 8], u[(c + q) % 8], x[c]);
    }
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < NCH; ++c) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace gdy
