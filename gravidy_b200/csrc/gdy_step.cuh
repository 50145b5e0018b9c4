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
    if (i < n) q[6 * JB] = s.m[i];
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
// Writes a2, a3, r, v, t, dt of particle i and returns the new step.
struct CorrIn { double dt1; double oa[3], oj[3], fa[3], fj[3]; Pred6 p; };
__device__ __forceinline__ double correct_core(const State &s, uint32_t i, double itime, double eta, const CorrIn &q)
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
    s.x[i] = add(add(q.p.x, mul(c4_24, a2[0])), mul(c5_120, a3[0]));
    s.y[i] = add(add(q.p.y, mul(c4_24, a2[1])), mul(c5_120, a3[1]));
    s.z[i] = add(add(q.p.z, mul(c4_24, a2[2])), mul(c5_120, a3[2]));
    s.vx[i] = add(add(q.p.vx, mul(c3_6, a2[0])), mul(c4_24, a3[0]));
    s.vy[i] = add(add(q.p.vy, mul(c3_6, a2[1])), mul(c4_24, a3[1]));
    s.vz[i] = add(add(q.p.vz, mul(c3_6, a2[2])), mul(c4_24, a3[2]));
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
    const double ndt = normalize_dt(normal_dt, dt1, itime);
    s.dt[i] = ndt;
    return ndt;
}

__global__ void k_correct(State s, uint32_t nact, double itime, double eta,
                          const double *__restrict__ fbuf, uint32_t ld)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nact) return;
    const uint32_t i = s.move[k];
    CorrIn q;
    q.dt1 = s.dt[i];
#pragma unroll
    for (int c = 0; c < 3; ++c)
    {
        if (fbuf)
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
k_min_final(const double *__restrict__ block_min, uint32_t nmin, double *atime_dev, double *atime_host)
{
    __shared__ double sm[8];
    const double m = fold_block_min(block_min, nmin, sm);
    if (threadIdx.x == 0)
    {
        *atime_dev = m;
        if (atime_host) *atime_host = m;
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
// The reference's triangular j>i loop is load-imbalanced; here every i sums over all j != i on a
// shared-memory tile and the result is halved.
__global__ void __launch_bounds__(128)
k_energy(State s, uint32_t n, uint32_t npad, double *__restrict__ ekin, double *__restrict__ epot)
{
    __shared__ double sx[128], sy[128], sz[128], smass[128];
    const uint32_t i = blockIdx.x * 128 + threadIdx.x;
    const uint32_t ic = i < n ? i : n - 1;
    const double xi = s.x[ic], yi = s.y[ic], zi = s.z[ic];
    double pot = 0.0;
    for (uint32_t t = 0; t < npad; t += 128)
    {
        __syncthreads();
        sx[threadIdx.x] = s.x[t + threadIdx.x];
        sy[threadIdx.x] = s.y[t + threadIdx.x];
        sz[threadIdx.x] = s.z[t + threadIdx.x];
        smass[threadIdx.x] = s.m[t + threadIdx.x];
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < 128; ++j)
        {
            const double dx = sx[j] - xi, dy = sy[j] - yi, dz = sz[j] - zi;
            const double r2 = fma(dz, dz, fma(dy, dy, dx * dx));
            double y = rsqrt_f64(r2);
            y = (r2 > 0.0) ? y : 0.0;   // i == j
            pot = fma(smass[j], y, pot);
        }
    }
    if (i < n)
    {
        const double mi = s.m[i];
        const double vx = s.vx[i], vy = s.vy[i], vz = s.vz[i];
        ekin[i] = 0.5 * mi * (vx * vx + vy * vy + vz * vz);
        epot[i] = -0.5 * mi * pot;
    }
}

// ---- k-th nearest neighbour distance ------------------------------------------------------------------
// The O(N^2) part of the Casertano & Hut density centre (reference: NbodyUtils::get_center_of_density,
// src/include/utils/NbodyUtils.cpp:163-220, sorts all N distances of every particle on the host).
// One thread per particle keeps the KNN_MAX smallest squared distances (self excluded) in registers,
// j through shared-memory tiles.  Padding particles sit 1e10 away and never enter the list.
constexpr int KNN_MAX = 8;
__global__ void __launch_bounds__(128)
k_knn_distance(State s, uint32_t n, uint32_t npad, int k, double *__restrict__ dist)
{
    __shared__ double sx[128], sy[128], sz[128];
    const uint32_t i = blockIdx.x * 128 + threadIdx.x;
    const uint32_t ic = i < n ? i : n - 1;
    const double xi = s.x[ic], yi = s.y[ic], zi = s.z[ic];
    double best[KNN_MAX];
#pragma unroll
    for (int q = 0; q < KNN_MAX; ++q) best[q] = 1e300;
    for (uint32_t t = 0; t < npad; t += 128)
    {
        __syncthreads();
        sx[threadIdx.x] = s.x[t + threadIdx.x];
        sy[threadIdx.x] = s.y[t + threadIdx.x];
        sz[threadIdx.x] = s.z[t + threadIdx.x];
        __syncthreads();
        for (int j = 0; j < 128; ++j)
        {
            const double dx = sx[j] - xi, dy = sy[j] - yi, dz = sz[j] - zi;
            double r2 = fma(dz, dz, fma(dy, dy, dx * dx));
            if (t + j == ic) continue;            // self
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
