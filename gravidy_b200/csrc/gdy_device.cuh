// gdy_device.cuh — device-side building blocks shared by every kernel of the Hermite4 hot path.
// sm_100a only.  No reference code is reproduced here; reference citations say what each piece
// replaces or must agree with.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gdy {

// Layout granules.
constexpr int JB = 64;           // particles per block of the blocked predicted array (one TMA bulk copy)
constexpr int JROWS = 7;         // rows of a block: x y z vx vy vz m
constexpr int JBD = JROWS * JB;  // doubles per block (3584 B)
constexpr int PAD = 256;         // arrays are padded to a multiple of this (zero-mass far particles)
constexpr double PAD_POS = 1.0e10;  // where the padding particles sit (mass 0 => contribute exactly 0)

// Device-resident particle state.  Plain arrays are structure-of-arrays, npad doubles each.
// Reference data contract: double4 h_r{x,y,z,m}, double4 h_v, Forces h_f/h_old, double3 h_a2/h_a3,
// double h_t/h_dt, Predictor h_p (reference: src/include/NbodySystem.hpp:100-115,
// src/include/common.hpp:213-231).  Mass is a float-representable double (Predictor.m is float).
//
// The predicted set (the reference's h_p / d_p, what every force evaluation streams) is kept BLOCKED:
// block b holds particles [64b, 64b+64) as 7 rows of 64 doubles (x y z vx vy vz m), 3584 contiguous
// bytes, so that one cp.async.bulk (TMA) brings a whole j-tile into shared memory, already in the
// row layout the inner loop reads with broadcast LDS.128.
struct State
{
    double *x, *y, *z, *m;          // h_r
    double *vx, *vy, *vz;           // h_v
    double *f[6];                   // h_f   : ax ay az jx jy jz
    double *old[6];                 // h_old
    double *a2[3], *a3[3];          // h_a2, h_a3
    double *t, *dt;                 // h_t, h_dt
    double *pb;                     // h_p, blocked: [npad/64][7][64]
    uint32_t *move;                 // h_move
    double *fsum;                   // [6][ld] summed partial forces of the active list (k-indexed)
    double *part;                   // [S][6][ld] j-split partial forces
    uint32_t *tile_counter;         // arrival counters of the fused last-block reduction
};

// index of row c of particle i in the blocked predicted array
__host__ __device__ __forceinline__ size_t pbidx(uint32_t i, int c)
{
    return (size_t)(i >> 6) * JBD + (size_t)c * JB + (i & 63);
}

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier + 1-D TMA bulk copy (cp.async.bulk, SASS UBLKCP) -----------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) { }
}
// global -> shared bulk copy, completion counted in bytes on `bar`.  bytes % 16 == 0, both
// addresses 16-byte aligned.
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                         uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// order this thread's (and, after a __syncwarp, its warp's) generic-proxy shared-memory reads before a
// following async-proxy (TMA) write to the same buffer
__device__ __forceinline__ void fence_proxy_async_smem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- FP64 reciprocal square root -------------------------------------------------------------
// Seed: rsqrt.approx.ftz.f64 (SASS MUFU.RSQ64H: the SFU's FP32-precision table applied to the high
// word, ~2^-22 relative error, runs on the XU pipe, not the FP64 pipe).  One cubically convergent
// correction y1 = y0 + y0*e*(1/2 + 3e/8), e = 1 - a*y0^2, leaves 5e^3/16 ~ 2^-64 of truncation error:
// full double precision for 5 FP64-pipe instructions.  Replaces rsqrt() in k_force_calculation
// (reference: src/include/gpu/Hermite4GPU_kernels.cu:101) / 1.0/sqrt() in the CPU oracle
// (reference: src/include/cpu/Hermite4CPU.cpp:66).
__device__ __forceinline__ double rsqrt_seed(double a)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
    return y;
}
__device__ __forceinline__ double rsqrt_f64(double a)
{
    const double y0 = rsqrt_seed(a);
    const double t = a * y0;
    const double e = fma(-t, y0, 1.0);
    const double p = fma(0.375, e, 0.5);
    const double ep = e * p;
    return fma(y0, ep, y0);   // two distinct register sources: a DFMA with three runs at 3/4 rate (tools/fp64_mix.cu)
}

// ---- the pair interaction -----------------------------------------------------------------------
// acc += m r / |r|^3 ; jerk += m ( v/|r|^3 - 3 (r.v) r / |r|^5 ), |r|^2 softened by e2.
// Same quantities as Hermite4CPU::force_calculation (reference: src/include/cpu/Hermite4CPU.cpp:55-82)
// arranged for the FP64 pipe: 6 DADD + 3 DFMA (r2) + DMUL,2 DFMA (r.v) + 5 (rsqrt) + 3 DMUL + 2 DMUL +
// 3 DFMA (acc) + 6 DFMA (jerk) = 31 FP64-pipe instructions per pair, FMA-contracted (tolerance of
// the north star for acc/jerk is 1e-12 relative; contraction and summation order are within it).
struct Acc6 { double ax, ay, az, jx, jy, jz; };

template <bool MASK_SELF>
__device__ __forceinline__ void pair(Acc6 &s, double xi, double yi, double zi, double vxi,
                                     double vyi, double vzi, double xj, double yj, double zj,
                                     double vxj, double vyj, double vzj, double mj, double e2)
{
    const double dx = xj - xi, dy = yj - yi, dz = zj - zi;
    const double dvx = vxj - vxi, dvy = vyj - vyi, dvz = vzj - vzi;
    double r2 = fma(dx, dx, e2);
    r2 = fma(dy, dy, r2);
    r2 = fma(dz, dz, r2);
    double rv = dx * dvx;
    rv = fma(dy, dvy, rv);
    rv = fma(dz, dvz, rv);
    double y = rsqrt_f64(r2);
    if (MASK_SELF) y = (r2 > 0.0) ? y : 0.0;   // e2 == 0: drop the i == j self pair (CPU oracle skips it)
    const double y2 = y * y;
    const double mr3 = (mj * y) * y2;
    const double c = (rv * y2) * (-3.0);
    s.ax = fma(mr3, dx, s.ax);
    s.ay = fma(mr3, dy, s.ay);
    s.az = fma(mr3, dz, s.az);
    s.jx = fma(mr3, fma(c, dx, dvx), s.jx);
    s.jy = fma(mr3, fma(c, dy, dvy), s.jy);
    s.jz = fma(mr3, fma(c, dz, dvz), s.jz);
}


// The same pair in two halves: pair_head leaves (d, dv, m/r^3, c) in registers, pair_tail does the nine
// three-source DFMAs.  Written apart so that the DFMAs that share m/r^3 (six) and c (three) can sit next to each
// other in the SASS: an operand the previous DFMA of the warp flagged `.reuse` is not read from the register file,
// which is what limits a three-source DFMA to 3/4 rate (tools/fp64_reuse.cu, profiles/archive/r01_fp64_reuse_v9.md).
struct PairMid { double dx, dy, dz, dvx, dvy, dvz, mr3, c; };

template <bool MASK_SELF>
__device__ __forceinline__ PairMid pair_head(double xi, double yi, double zi, double vxi, double vyi, double vzi,
                                             double xj, double yj, double zj, double vxj, double vyj, double vzj,
                                             double mj, double e2)
{
    PairMid p;
    p.dx = xj - xi; p.dy = yj - yi; p.dz = zj - zi;
    p.dvx = vxj - vxi; p.dvy = vyj - vyi; p.dvz = vzj - vzi;
    double r2 = fma(p.dx, p.dx, e2);
    r2 = fma(p.dy, p.dy, r2);
    r2 = fma(p.dz, p.dz, r2);
    double rv = p.dx * p.dvx;
    rv = fma(p.dy, p.dvy, rv);
    rv = fma(p.dz, p.dvz, rv);
    double y = rsqrt_f64(r2);
    if (MASK_SELF) y = (r2 > 0.0) ? y : 0.0;
    const double y2 = y * y;
    p.mr3 = (mj * y) * y2;
    p.c = (rv * y2) * (-3.0);
    return p;
}

__device__ __forceinline__ void pair_tail(Acc6 &s, const PairMid &p)
{
    const double tx = fma(p.c, p.dx, p.dvx);
    const double ty = fma(p.c, p.dy, p.dvy);
    const double tz = fma(p.c, p.dz, p.dvz);
    s.ax = fma(p.mr3, p.dx, s.ax);
    s.ay = fma(p.mr3, p.dy, s.ay);
    s.az = fma(p.mr3, p.dz, s.az);
    s.jx = fma(p.mr3, tx, s.jx);
    s.jy = fma(p.mr3, ty, s.jy);
    s.jz = fma(p.mr3, tz, s.jz);
}

// Software-pipelined halves (GDY_FORCE_FORM 3): pair_head7 also does t = c*d + dv, so that what is carried to the
// next trip is (d, t, m/r^3) and the tail is six DFMAs that all share m/r^3.  A zeroed PairCarry adds exactly 0.
struct PairCarry { double dx, dy, dz, tx, ty, tz, mr3; };

template <bool MASK_SELF>
__device__ __forceinline__ PairCarry pair_head7(double xi, double yi, double zi, double vxi, double vyi, double vzi,
                                                double xj, double yj, double zj, double vxj, double vyj, double vzj,
                                                double mj, double e2)
{
    const PairMid p = pair_head<MASK_SELF>(xi, yi, zi, vxi, vyi, vzi, xj, yj, zj, vxj, vyj, vzj, mj, e2);
    PairCarry c;
    c.dx = p.dx; c.dy = p.dy; c.dz = p.dz;
    c.tx = fma(p.c, p.dx, p.dvx);
    c.ty = fma(p.c, p.dy, p.dvy);
    c.tz = fma(p.c, p.dz, p.dvz);
    c.mr3 = p.mr3;
    return c;
}

// the second half of pair_head7 on its own (form 6 keeps the three DFMAs that share c in a block of their own)
__device__ __forceinline__ PairCarry pair_carry(const PairMid &p)
{
    PairCarry c;
    c.dx = p.dx; c.dy = p.dy; c.dz = p.dz;
    c.tx = fma(p.c, p.dx, p.dvx);
    c.ty = fma(p.c, p.dy, p.dvy);
    c.tz = fma(p.c, p.dz, p.dvz);
    c.mr3 = p.mr3;
    return c;
}

__device__ __forceinline__ void pair_tail6(Acc6 &s, const PairCarry &p)
{
    s.ax = fma(p.mr3, p.dx, s.ax);
    s.ay = fma(p.mr3, p.dy, s.ay);
    s.az = fma(p.mr3, p.dz, s.az);
    s.jx = fma(p.mr3, p.tx, s.jx);
    s.jy = fma(p.mr3, p.ty, s.jy);
    s.jz = fma(p.mr3, p.tz, s.jz);
}

}  // namespace gdy
