// gdy_force.cuh — the force kernels: FP64 acc+jerk of the active i-particles from a j-range.
//
// Replaces k_init_acc_jrk, k_update and k_reduce (reference: src/include/gpu/Hermite4GPU_kernels.cu:
// 47-87, 165-219, 226-248).  The reference uses 32-thread CTAs, an AoS j-tile of 32 in shared memory, a
// fixed 16-way j-split and one thread per i whatever nact is.  Here:
//
//   k_force_warp  (large and medium blocks): i-particles in registers (IPT per lane, 32 lanes), j-tiles
//       of 64 particles = one 3584-byte block of the blocked predicted array, brought into a PER-WARP
//       shared-memory ring by one cp.async.bulk (TMA, SASS UBLKCP) per tile with mbarrier complete_tx;
//       every lane reads the same j (broadcast LDS.128 = two j per load).  Warps never wait for each
//       other inside the j-loop (no CTA barrier per tile): a warp is a self-contained pipeline.  The
//       warps of a CTA are arranged WI x WJ: WI different i-groups, WJ interleaved j-subranges of the
//       same i-group, folded through shared memory at the end.
//   k_force_lane  (small blocks): the warp is IL i-lanes x JL = 32/IL j-lanes; every j-lane streams its
//       own j-particles straight from L2 (coalesced, no staging: each j is used IL*IPT times only),
//       accumulators are folded over the j-lanes with xor-shuffles and over the warps through shared
//       memory.  A block step with nact = 2 still spreads the N j-particles over every SM.  The j-split
//       partials are summed by the LAST CTA to finish (arrival counter), in fixed split order, inside
//       the same launch.
//
// Both write j-split partial sums part[split][6][ld]; the split dimension is always summed in a fixed
// order (deterministic, so replicated multi-GPU state stays bit-identical).
#pragma once
#include "gdy_device.cuh"

// Inner-loop form of k_force_warp.  0 = one pair() after the other; 1 = heads of the 2 x IPT pairs of a trip, then
// their tails (pair_head / pair_tail; putting a tail into ONE asm statement changes nothing: ptxas reorders it);
// 3 = software-pipelined: the six-DFMA tails of the previous trip's pairs, which all share m/r^3, issue while this
// trip's LDS.128 are in flight, then this trip's heads.  tools/force_lab.sh builds them side by side;
// measured (profiles/archive/r01_force_lab_v9.md): form 3, unroll 1, 512 resident threads is the fastest, 508 Gpairs/s.
// 6 = form 3 with the instruction groups PINNED (IPT = 2 only, the other shapes keep form 3): every group of FP64
// instructions that shares an operand — the six-DFMA tail of a pair (m/r^3), the three t = c*d + dv of a pair (c) — sits
// in a basic block of its own, so ptxas cannot interleave the groups and flags the shared register `.reuse` from one
// DFMA to the next (a three-source DFMA fed by the operand reuse cache costs 2.0 instead of 2.96 cycles, section 3.1
// of DESIGN.md).  The block boundaries are labels that a never-taken `goto` ahead of the loop also jumps to (the
// condition is bits of ld / per_split that are never set, opaque to the compiler): a join point is the one thing
// ptxas does not schedule across, and it costs no instruction.  Same arithmetic in the same order per accumulator
// as form 3: bit-identical results.  SASS of the loop: 28 of the 44 three-source DFMAs of a trip fed (form 3: 13),
// modelled 65.8 instead of 69.4 cycles per pair (tools/sass_fp64_cost.py; profiles/r02_force_lab.md).
#ifndef GDY_FORCE_FORM
#define GDY_FORCE_FORM 6
#endif
#define GDY_FORM_PIPE (GDY_FORCE_FORM == 3 || GDY_FORCE_FORM == 6)
#ifndef GDY_FORCE_UNROLL
#define GDY_FORCE_UNROLL 1
#endif
// threads the warp kernels are compiled to keep resident per SM (register cap = 65536 / this)
#ifndef GDY_FORCE_RESIDENT
#define GDY_FORCE_RESIDENT 512
#endif
#define GDY_STR_(x) #x
#define GDY_UNROLL(n) _Pragma(GDY_STR_(unroll n))

namespace gdy {

// Arguments common to every force kernel (one struct: one signature, one launch table).
struct ForceArgs
{
    const double *pb;          // blocked predicted j-particles (and the i-particles, gathered by index)
    const uint32_t *move;      // active list (k -> particle index); NULL = identity (full-active sweep)
    double *part;              // [nsplit][6][ld]
    double *fsum;              // [6][ld]   (lane kernels: written by the last CTA)
    uint32_t *counter;         // [grid.x]  (lane kernels: arrival counters, left at 0)
    uint32_t nact, ld;
    uint32_t jb, je;           // j-range of this shard: blocks of 64 (warp kernels) / particles (lane kernels)
    uint32_t per_split;        // same unit
    double e2;
};

struct IReg { double x, y, z, vx, vy, vz; };

// two consecutive doubles from shared memory at byte address sa + OFF: LDS.128 with an immediate offset, so that the
// seven rows of a j pair hang off ONE address register (every instruction that reads a register competes with the
// FP64 pipe's operand reads: profiles/r02_force_lab.md)
template <int OFF>
__device__ __forceinline__ double2 lds2(uint32_t sa)
{
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(v.x), "=d"(v.y) : "r"(sa), "n"(OFF) : "memory");
    return v;
}

__device__ __forceinline__ IReg load_i(const ForceArgs &a, uint32_t k)
{
    const uint32_t kc = k < a.nact ? k : a.nact - 1;   // lanes past nact compute a duplicate, never stored
    const uint32_t i = a.move ? a.move[kc] : kc;
    const double *p = a.pb + pbidx(i, 0);
    return IReg{p[0], p[JB], p[2 * JB], p[3 * JB], p[4 * JB], p[5 * JB]};
}

// ------------------------------------------------------------------------------------------------------
template <int IPT, int WI, int WJ, int NST, bool MASK>
__global__ void __launch_bounds__(32 * WI * WJ, GDY_FORCE_RESIDENT / (32 * WI * WJ))
k_force_warp(const ForceArgs a)
{
    constexpr int NW = WI * WJ;
    constexpr int TI = 32 * WI * IPT;   // i-particles per CTA
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *ring = reinterpret_cast<double *>(smem_raw);                          // [NW][NST][JBD]
    uint64_t *full = reinterpret_cast<uint64_t *>(ring + (size_t)NW * NST * JBD);  // [NW][NST]

    const int tid = threadIdx.x, lane = tid & 31;
    // The warp index through a warp reduction (REDUX writes a uniform register): ptxas then KNOWS that it — and the ring
    // address, the tile count and the loop's address arithmetic derived from it — is the same in every lane, keeps them in
    // uniform registers, and the hot loop's LDS.128 [UR + imm] / UIADD3 / UISETP / BRA.U stop reading vector registers.
    // Every instruction that reads a vector register takes an operand-read cycle away from the FP64 pipe: nine of
    // them per trip gone, 552 instead of 538 Gpairs/s (profiles/r02_force_lab.md).
    const int warp = (int)__reduce_max_sync(0xffffffffu, (unsigned)(tid >> 5));
    const int wi = warp % WI, wj = warp / WI;
    double *myring = ring + (size_t)warp * NST * JBD;
    uint64_t *myfull = full + warp * NST;

    // this CTA's j-blocks, then this warp's (interleaved over the WJ j-warps)
    const uint32_t b0 = a.jb + blockIdx.y * a.per_split;
    const uint32_t b1 = min(b0 + a.per_split, a.je);
    const int nb_cta = b1 > b0 ? (int)(b1 - b0) : 0;
    const int nt = nb_cta > wj ? (nb_cta - wj + WJ - 1) / WJ : 0;   // tiles of this warp

    if (lane == 0)
    {
#pragma unroll
        for (int s = 0; s < NST; ++s) mbar_init(&myfull[s], 1);
        mbar_fence_init();
    }
    __syncthreads();

    auto issue = [&](int t) {
        const int s = t % NST;
        const double *src = a.pb + (size_t)(b0 + wj + (uint32_t)t * WJ) * JBD;
        mbar_expect_tx(&myfull[s], JBD * sizeof(double));
        bulk_g2s(myring + (size_t)s * JBD, src, JBD * sizeof(double), &myfull[s]);
    };
    if (lane == 0)
    {
        const int pre = nt < NST ? nt : NST;
        for (int t = 0; t < pre; ++t) issue(t);
    }

    uint32_t k[IPT];
    IReg ri[IPT];
    Acc6 acc[IPT];
#pragma unroll
    for (int q = 0; q < IPT; ++q)
    {
        k[q] = blockIdx.x * TI + (wi * IPT + q) * 32 + lane;
        ri[q] = load_i(a, k[q]);
        acc[q] = Acc6{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    }
    const double e2 = a.e2;
#if GDY_FORM_PIPE
    PairCarry pc[2 * IPT];
#pragma unroll
    for (int q = 0; q < 2 * IPT; ++q) pc[q] = PairCarry{0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#endif

#if GDY_FORCE_FORM == 6
    if constexpr (IPT == 2)
    {
        // everything the never-taken entries could reach is initialised: no undefined value on any path
        double2 X = make_double2(0.0, 0.0), Y = X, Z = X, VX = X, VY = X, VZ = X, M = X;
        PairMid pm[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) pm[q] = PairMid{0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        // the loop's one induction variable: the shared-memory byte address of this trip's two j (the seven rows sit at
        // immediate offsets from it), compared with the end of the tile's first row
        const uint32_t ring_u = smem_u32(myring);   // warp-uniform (see `warp` above)
        uint32_t sa = ring_u, send = sa + JB * 8;
        int t = 0;
        // never set (ld, nact <= 2^28, per_split, jb, je <= 2^24): the entries below exist to make every label a join point
        const uint32_t opq = (a.ld >> 29) | ((a.per_split >> 25) << 3) | ((a.nact >> 29) << 10) | ((a.jb >> 26) << 13) |
                             ((a.je >> 26) << 19);
        if (nt > 0)
        {
            mbar_wait(&myfull[0], 0u);
            if (opq & 0x001u) goto G_LX;
            if (opq & 0x002u) goto G_T1;
            if (opq & 0x004u) goto G_LV;
            if (opq & 0x008u) goto G_T2;
            if (opq & 0x010u) goto G_LM;
            if (opq & 0x020u) goto G_T3;
            if (opq & 0x040u) goto G_HEAD;
            if (opq & 0x080u) goto G_C0;
            if (opq & 0x100u) goto G_C1;
            if (opq & 0x200u) goto G_C2;
            if (opq & 0x400u) goto G_C3;
            for (;;)
            {
                // tails of the previous trip's four pairs, the loads of this trip's j pair dealt out between them (a
                // tail frees seven doubles of carry, the next loads land in them: no register pressure peak)
                pair_tail6(acc[0], pc[0]);
            G_LX:
                X = lds2<0 * JB * 8>(sa);
                Y = lds2<1 * JB * 8>(sa);
                Z = lds2<2 * JB * 8>(sa);
            G_T1:
                pair_tail6(acc[1], pc[1]);
            G_LV:
                VX = lds2<3 * JB * 8>(sa);
                VY = lds2<4 * JB * 8>(sa);
                VZ = lds2<5 * JB * 8>(sa);
            G_T2:
                pair_tail6(acc[0], pc[2]);
            G_LM:
                M = lds2<6 * JB * 8>(sa);
            G_T3:
                pair_tail6(acc[1], pc[3]);
            G_HEAD:
                // heads: differences, r^2, r.v, rsqrt, m/r^3, c — four independent chains for ptxas to interleave
#pragma unroll
                for (int q = 0; q < 2; ++q)
                    pm[q] = pair_head<MASK>(ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, X.x, Y.x, Z.x,
                                            VX.x, VY.x, VZ.x, M.x, e2);
#pragma unroll
                for (int q = 0; q < 2; ++q)
                    pm[2 + q] = pair_head<MASK>(ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, X.y, Y.y, Z.y,
                                                VX.y, VY.y, VZ.y, M.y, e2);
            G_C0:
                pc[0] = pair_carry(pm[0]);   // t = c*d + dv: three DFMAs that share c
            G_C1:
                pc[1] = pair_carry(pm[1]);
            G_C2:
                pc[2] = pair_carry(pm[2]);
            G_C3:
                pc[3] = pair_carry(pm[3]);
                sa += 16;
                if (sa != send) continue;
                __syncwarp();   // every lane is done with this stage -> it may be refilled
                if (lane == 0 && t + NST < nt)
                {
                    fence_proxy_async_smem();
                    issue(t + NST);
                }
                if (++t >= nt) break;
                const int s = t % NST;
                mbar_wait(&myfull[s], (uint32_t)(t / NST) & 1u);
                sa = ring_u + (uint32_t)s * (JBD * 8);
                send = sa + JB * 8;
            }
            // The last trip's tails, HERE and not with the common drain below: nvcc 12.9 drops that drain for this
            // branch (it is gone from the PTX when the values come out of the multi-entry loop; found by
            // tools/lab/force_probe.py: the last two j of every warp's last tile contributed nothing).  The carries are
            // zeroed so that the common drain adds exactly 0.
            pair_tail6(acc[0], pc[0]); pair_tail6(acc[1], pc[1]); pair_tail6(acc[0], pc[2]); pair_tail6(acc[1], pc[3]);
#pragma unroll
            for (int q = 0; q < 4; ++q) pc[q] = PairCarry{0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        }
    }
    else
#endif
    for (int t = 0; t < nt; ++t)
    {
        const int s = t % NST;
        mbar_wait(&myfull[s], (uint32_t)(t / NST) & 1u);
        const double *b = myring + (size_t)s * JBD;
        GDY_UNROLL(GDY_FORCE_UNROLL)
        for (int jj = 0; jj < JB; jj += 2)
        {
            const double2 X = *reinterpret_cast<const double2 *>(b + 0 * JB + jj);
            const double2 Y = *reinterpret_cast<const double2 *>(b + 1 * JB + jj);
            const double2 Z = *reinterpret_cast<const double2 *>(b + 2 * JB + jj);
            const double2 VX = *reinterpret_cast<const double2 *>(b + 3 * JB + jj);
            const double2 VY = *reinterpret_cast<const double2 *>(b + 4 * JB + jj);
            const double2 VZ = *reinterpret_cast<const double2 *>(b + 5 * JB + jj);
            const double2 M = *reinterpret_cast<const double2 *>(b + 6 * JB + jj);
#if GDY_FORM_PIPE
            // software pipeline: the six-DFMA tails of the previous trip's pairs issue while this trip's LDS.128
            // are in flight, then this trip's heads
#pragma unroll
            for (int q = 0; q < 2 * IPT; ++q) pair_tail6(acc[q % IPT], pc[q]);
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pc[q] = pair_head7<MASK>(ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, X.x, Y.x, Z.x,
                                         VX.x, VY.x, VZ.x, M.x, e2);
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pc[IPT + q] = pair_head7<MASK>(ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, X.y, Y.y,
                                               Z.y, VX.y, VY.y, VZ.y, M.y, e2);
#elif GDY_FORCE_FORM == 1
            // heads of the 2 x IPT pairs of this trip, then their nine-DFMA tails
            PairMid pm[2 * IPT];
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pm[q] = pair_head<MASK>(ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, X.x, Y.x, Z.x,
                                        VX.x, VY.x, VZ.x, M.x, e2);
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pm[IPT + q] = pair_head<MASK>(ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, X.y, Y.y,
                                              Z.y, VX.y, VY.y, VZ.y, M.y, e2);
#pragma unroll
            for (int q = 0; q < 2 * IPT; ++q) pair_tail(acc[q % IPT], pm[q]);
#else
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pair<MASK>(acc[q], ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, X.x, Y.x, Z.x,
                           VX.x, VY.x, VZ.x, M.x, e2);
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pair<MASK>(acc[q], ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, X.y, Y.y, Z.y,
                           VX.y, VY.y, VZ.y, M.y, e2);
#endif
        }
        __syncwarp();   // every lane is done with stage s -> it may be refilled
        if (lane == 0 && t + NST < nt)
        {
            fence_proxy_async_smem();
            issue(t + NST);
        }
    }

#if GDY_FORM_PIPE
#pragma unroll
    for (int q = 0; q < 2 * IPT; ++q) pair_tail6(acc[q % IPT], pc[q]);
#endif

    double *out = a.part + (size_t)blockIdx.y * 6 * a.ld;
    if (WJ == 1)
    {
#pragma unroll
        for (int q = 0; q < IPT; ++q)
            if (k[q] < a.nact)
            {
                out[0 * (size_t)a.ld + k[q]] = acc[q].ax;
                out[1 * (size_t)a.ld + k[q]] = acc[q].ay;
                out[2 * (size_t)a.ld + k[q]] = acc[q].az;
                out[3 * (size_t)a.ld + k[q]] = acc[q].jx;
                out[4 * (size_t)a.ld + k[q]] = acc[q].jy;
                out[5 * (size_t)a.ld + k[q]] = acc[q].jz;
            }
    }
    else
    {
        // fold the WJ j-warps in fixed order: red[wj][wi][q][c][lane], aliased on the (now idle) rings
        static_assert((size_t)WJ * WI * IPT * 6 * 32 <= (size_t)NW * NST * JBD, "reduce buffer must fit the rings");
        __syncthreads();
        double *red = ring;
#pragma unroll
        for (int q = 0; q < IPT; ++q)
        {
            double *r = red + ((size_t)((wj * WI + wi) * IPT + q) * 6) * 32 + lane;
            r[0 * 32] = acc[q].ax; r[1 * 32] = acc[q].ay; r[2 * 32] = acc[q].az;
            r[3 * 32] = acc[q].jx; r[4 * 32] = acc[q].jy; r[5 * 32] = acc[q].jz;
        }
        __syncthreads();
        if (wj == 0)
        {
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                if (k[q] < a.nact)
                {
#pragma unroll
                    for (int c = 0; c < 6; ++c)
                    {
                        double sum = 0.0;
#pragma unroll
                        for (int w = 0; w < WJ; ++w)
                            sum += red[((size_t)((w * WI + wi) * IPT + q) * 6 + c) * 32 + lane];
                        out[c * (size_t)a.ld + k[q]] = sum;
                    }
                }
        }
    }
}

template <int IPT, int WI, int WJ, int NST>
constexpr size_t force_warp_smem()
{
    return (size_t)WI * WJ * NST * JBD * sizeof(double) + (size_t)WI * WJ * NST * sizeof(uint64_t);
}

// ------------------------------------------------------------------------------------------------------
template <int IL, int IPT, int WJ, bool MASK>
__global__ void __launch_bounds__(32 * WJ)
k_force_lane(const ForceArgs a)
{
    constexpr int JL = 32 / IL;       // j-lanes per warp
    constexpr int TI = IL * IPT;      // i-particles per CTA
    constexpr int PJ = JL * WJ;       // j-particles in flight per CTA step
    constexpr int U = 4;              // independent j per lane per trip (memory-level parallelism)
    constexpr int NT = 32 * WJ;       // threads
    constexpr int NO = IPT * 6 * IL;  // outputs of this CTA
    constexpr int G = NT / NO > 0 ? NT / NO : 1;   // split-groups of the final reduction

    __shared__ double red[WJ][IPT][6][IL];
    __shared__ double red2[G][NO];
    __shared__ uint32_t s_last;

    const int tid = threadIdx.x, lane = tid & 31, wj = tid >> 5;
    const int il = lane % IL, jl = lane / IL;

    const uint32_t jb = a.jb + blockIdx.y * a.per_split;
    const uint32_t je = min(jb + a.per_split, a.je);

    uint32_t k[IPT];
    IReg ri[IPT];
    Acc6 acc[IPT];
#pragma unroll
    for (int q = 0; q < IPT; ++q)
    {
        k[q] = blockIdx.x * TI + q * IL + il;
        ri[q] = load_i(a, k[q]);
        acc[q] = Acc6{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    }
    const double e2 = a.e2;

    for (uint32_t base = jb; base < je; base += PJ * U)
    {
        double xj[U], yj[U], zj[U], vxj[U], vyj[U], vzj[U], mj[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const uint32_t j = base + (u * WJ + wj) * JL + jl;
            const bool ok = j < je;
            const double *p = a.pb + pbidx(ok ? j : jb, 0);
            xj[u] = __ldg(p); yj[u] = __ldg(p + JB); zj[u] = __ldg(p + 2 * JB);
            vxj[u] = __ldg(p + 3 * JB); vyj[u] = __ldg(p + 4 * JB); vzj[u] = __ldg(p + 5 * JB);
            const double mm = __ldg(p + 6 * JB);
            mj[u] = ok ? mm : 0.0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pair<MASK>(acc[q], ri[q].x, ri[q].y, ri[q].z, ri[q].vx, ri[q].vy, ri[q].vz, xj[u], yj[u],
                           zj[u], vxj[u], vyj[u], vzj[u], mj[u], e2);
    }

    // fold the JL j-lanes (lane = jl*IL + il): xor offsets IL, 2IL, ..., 16
#pragma unroll
    for (int q = 0; q < IPT; ++q)
    {
#pragma unroll
        for (int off = IL; off < 32; off <<= 1)
        {
            acc[q].ax += __shfl_xor_sync(0xffffffffu, acc[q].ax, off);
            acc[q].ay += __shfl_xor_sync(0xffffffffu, acc[q].ay, off);
            acc[q].az += __shfl_xor_sync(0xffffffffu, acc[q].az, off);
            acc[q].jx += __shfl_xor_sync(0xffffffffu, acc[q].jx, off);
            acc[q].jy += __shfl_xor_sync(0xffffffffu, acc[q].jy, off);
            acc[q].jz += __shfl_xor_sync(0xffffffffu, acc[q].jz, off);
        }
        if (jl == 0)
        {
            red[wj][q][0][il] = acc[q].ax; red[wj][q][1][il] = acc[q].ay; red[wj][q][2][il] = acc[q].az;
            red[wj][q][3][il] = acc[q].jx; red[wj][q][4][il] = acc[q].jy; red[wj][q][5][il] = acc[q].jz;
        }
    }
    __syncthreads();
    // fold the WJ warps in fixed order; NO outputs per CTA.  Output o = (q, c, l).
    const uint32_t nsplit = gridDim.y;
    double *out = (nsplit == 1 ? a.fsum : a.part + (size_t)blockIdx.y * 6 * a.ld);
    for (int o = tid; o < NO; o += NT)
    {
        const int l = o % IL, c = (o / IL) % 6, q = o / (6 * IL);
        const uint32_t kk = blockIdx.x * TI + q * IL + l;
        if (kk < a.nact)
        {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < WJ; ++w) s += red[w][q][c][l];
            out[c * (size_t)a.ld + kk] = s;
        }
    }
    if (nsplit == 1) return;

    // ---- fused j-split reduction: the last CTA of this i-tile to arrive sums all splits ------------------
    __threadfence();
    __syncthreads();
    if (tid == 0)
    {
        const uint32_t prev = atomicAdd(&a.counter[blockIdx.x], 1u);
        s_last = (prev == nsplit - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // thread (g, o): splits g, g+G, ... in ascending order; then the G group sums in ascending g.
    // The decomposition depends only on the launch shape, never on which CTA happens to be last.
    for (int idx = tid; idx < G * NO; idx += NT)
    {
        const int o = idx % NO, g = idx / NO;
        const int l = o % IL, c = (o / IL) % 6, q = o / (6 * IL);
        const uint32_t kk = blockIdx.x * TI + q * IL + l;
        double s = 0.0;
        if (kk < a.nact)
        {
            const volatile double *p = a.part + c * (size_t)a.ld + kk;
            const size_t stride = (size_t)6 * a.ld;
            uint32_t sp = g;
            for (; sp + 7 * G < nsplit; sp += 8 * G)
            {
                double v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = p[(size_t)(sp + u * G) * stride];
#pragma unroll
                for (int u = 0; u < 8; ++u) s += v[u];
            }
            for (; sp < nsplit; sp += G) s += p[(size_t)sp * stride];
        }
        red2[g][o] = s;
    }
    __syncthreads();
    for (int o = tid; o < NO; o += NT)
    {
        const int l = o % IL, c = (o / IL) % 6, q = o / (6 * IL);
        const uint32_t kk = blockIdx.x * TI + q * IL + l;
        if (kk < a.nact)
        {
            double s = 0.0;
#pragma unroll
            for (int g = 0; g < G; ++g) s += red2[g][o];
            a.fsum[c * (size_t)a.ld + kk] = s;
        }
    }
    if (tid == 0) a.counter[blockIdx.x] = 0;   // ready for the next launch (stream-ordered)
}

// Sum the j-split partials of the warp kernels in ascending split order (deterministic) -> fsum[6][ld].
// Replaces k_reduce (reference: src/include/gpu/Hermite4GPU_kernels.cu:226-248).
// Block = 32 k-lanes x 8 split-warps: coalesced over k, the splits strided over the warps (4 loads in
// flight per thread) and folded through shared memory in warp order.
// With `xs.world > 1` the sums also leave for the other GPUs from here: flagged slots into buffer [parity][my rank] of
// every rank's exchange region over NVLink (the send half of the one-shot exchange, gdy_step.cuh), so that the reduce
// and the collective are one launch.
struct XSend
{
    uint4 *peer[8];                       // every rank's exchange buffer as seen from this device (XMID_MAXW)
    int rank, world;                      // world <= 1: nothing is sent
    uint32_t tag, par;
    uint32_t xld;                         // slots per (rank, component) row of the exchange region (XMID_LD)
};

__device__ __forceinline__ void xll_store_sys(uint4 *slot, double v, uint32_t tag)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    asm volatile("st.relaxed.sys.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(slot), "r"((uint32_t)b), "r"(tag),
                 "r"((uint32_t)(b >> 32)), "r"(tag)
                 : "memory");
}

__global__ void __launch_bounds__(256)
k_reduce_partials(const double *__restrict__ part, double *__restrict__ fsum, uint32_t nact,
                  uint32_t ld, uint32_t nsplit, const XSend xs)
{
    // grid (k-blocks of 32, 6 components): one component per CTA, so that a mid-size block with many splits (1172
    // active particles x 171 splits) is 222 CTAs with two rounds of loads each instead of 37 CTAs with a dozen.  Per
    // (k, component) the additions happen in the same order as ever: the splits w, w + 8, ... in ascending order in
    // warp w, then the eight warps in ascending order.
    __shared__ double red[8][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t k = blockIdx.x * 32 + lane, c = blockIdx.y;
    double acc = 0.0;
    if (k < nact)
    {
        const size_t stride = (size_t)6 * ld;
        const double *p = part + (size_t)c * ld + k;
        uint32_t sp = w;
        for (; sp + 56 < nsplit; sp += 64)
        {
            double v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = p[(size_t)(sp + 8 * u) * stride];
#pragma unroll
            for (int u = 0; u < 8; ++u) acc += v[u];
        }
        for (; sp < nsplit; sp += 8) acc += p[(size_t)sp * stride];
    }
    red[w][lane] = acc;
    __syncthreads();
    if (w == 0 && k < nact)
    {
        double sum = 0.0;
#pragma unroll
        for (int ww = 0; ww < 8; ++ww) sum += red[ww][lane];
        fsum[(size_t)c * ld + k] = sum;
        if (xs.world > 1)
        {
            const size_t at = (((size_t)xs.par * 8 + xs.rank) * 6 + c) * xs.xld + k;
            for (int r = 0; r < xs.world; ++r) xll_store_sys(xs.peer[r] + at, sum, xs.tag);
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Variant table.  kind 0 = warp kernel (j staged through shared memory by TMA), 1 = lane (j from L2).
using ForceFn = void (*)(const ForceArgs);

struct ForceVariant
{
    int kind;       // 0 warp, 1 lane
    int ti;         // i-particles per CTA
    int threads;    // CTA size
    int pj;         // lane kernels: j-particles per CTA trip (PJ*U); warp kernels: JB
    size_t smem;    // dynamic shared memory
    const char *name;
    ForceFn fn[2];  // [MASK]
};

enum : int
{
    FV_WARP_I256 = 0,   // IPT=2 WI=4 WJ=1
    FV_WARP_I128,       // IPT=2 WI=2 WJ=2
    FV_WARP_I64,        // IPT=2 WI=1 WJ=4
    FV_WARP_I32,        // IPT=1 WI=1 WJ=4
    FV_LANE_I16,        // IL=16
    FV_LANE_I8,         // IL=8
    FV_LANE_I4,         // IL=4
    FV_LANE_I2,         // IL=2
    FV_LANE_I1,         // IL=1
    FV_LANE_I1x2,       // IL=1 IPT=2
    FV_LANE_I1x4,       // IL=1 IPT=4
    FV_WARP_I512,       // IPT=2 WI=8 WJ=1
    FV_WARP_I128H,      // IPT=1 WI=4 WJ=1 (high occupancy)
    FV_LANE_I16x2,      // IL=16 IPT=2 (32 i per CTA)
    FV_LANE_I16x4,      // IL=16 IPT=4 (64 i per CTA)
    FV_LANE_I1x8,       // IL=1 IPT=8
    FV_COUNT
};

#define GDY_WARP(ipt, wi, wj, nst, nm) \
    {0, 32 * (wi) * (ipt), 32 * (wi) * (wj), JB, force_warp_smem<ipt, wi, wj, nst>(), nm, \
     {k_force_warp<ipt, wi, wj, nst, false>, k_force_warp<ipt, wi, wj, nst, true>}}
#define GDY_LANE(il, ipt, nm) \
    {1, (il) * (ipt), 128, (32 / (il)) * 4 * 4, 0, nm, \
     {k_force_lane<il, ipt, 4, false>, k_force_lane<il, ipt, 4, true>}}

static const ForceVariant g_variants[FV_COUNT] = {
    GDY_WARP(2, 4, 1, 3, "warp_i256"), GDY_WARP(2, 2, 2, 3, "warp_i128"), GDY_WARP(2, 1, 4, 3, "warp_i64"),
    GDY_WARP(1, 1, 4, 2, "warp_i32"),
    GDY_LANE(16, 1, "lane_i16"), GDY_LANE(8, 1, "lane_i8"), GDY_LANE(4, 1, "lane_i4"),
    GDY_LANE(2, 1, "lane_i2"), GDY_LANE(1, 1, "lane_i1"), GDY_LANE(1, 2, "lane_i1x2"),
    GDY_LANE(1, 4, "lane_i1x4"),
    GDY_WARP(2, 8, 1, 3, "warp_i512"), GDY_WARP(1, 4, 1, 2, "warp_i128h"),
    GDY_LANE(16, 2, "lane_i16x2"), GDY_LANE(16, 4, "lane_i16x4"), GDY_LANE(1, 8, "lane_i1x8"),
};

// opt in to > 48 KiB of dynamic shared memory where a variant needs it (once per process and device)
static cudaError_t force_variants_prepare()
{
    for (int v = 0; v < FV_COUNT; ++v)
        for (int m = 0; m < 2; ++m)
            if (g_variants[v].smem > 48 * 1024)
            {
                cudaError_t e = cudaFuncSetAttribute((const void *)g_variants[v].fn[m],
                                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                     (int)g_variants[v].smem);
                if (e != cudaSuccess) return e;
            }
    return cudaSuccess;
}

}  // namespace gdy
