// gdy_force.cuh — the force kernels: FP64 acc+jerk of the active i-particles from a j-range.
//
// Replaces k_init_acc_jrk, k_update and (together with k_reduce_partials in gdy_step.cuh) k_reduce
// (reference: src/include/gpu/Hermite4GPU_kernels.cu:47-87, 165-219, 226-248).  The reference uses
// 32-thread CTAs, an AoS j-tile of 32 in shared memory, a fixed 16-way j-split and one thread per i
// whatever nact is.  Here:
//
//   k_force_tile  (nact > 16): i-particles in registers (IPT per thread), 32 i per warp; j-tiles of
//       TJ=128 particles x 7 SoA arrays are staged into a 4-deep shared-memory ring by 1-D TMA bulk
//       copies (cp.async.bulk + mbarrier complete_tx) issued by one elected thread; every lane reads
//       the same j (broadcast LDS.128 = two j per load).  WJ > 1 lets the warps of a CTA share one
//       32-wide i-tile and split the j-tile, then sum through shared memory.
//   k_force_lane  (nact <= 16): the warp is IL i-lanes x JL = 32/IL j-lanes; every j-lane streams its
//       own j-particles straight from global/L2 (coalesced, no staging: each j is used IL*IPT times
//       only), accumulators are folded over the j-lanes with xor-shuffles and over the warps through
//       shared memory.  This is the "small active block" path: a block step with nact = 2 still
//       spreads N j-particles over every SM.
//
// Both write j-split partial sums part[split][6][nact_pad]; the split dimension is summed in a fixed
// order afterwards (deterministic, so replicated multi-GPU state stays bit-identical).
#pragma once
#include "gdy_device.cuh"

namespace gdy {

// ------------------------------------------------------------------------------------------------------
template <int IPT, int WI, int WJ, bool MASK>
__global__ void __launch_bounds__(32 * WI * WJ)
k_force_tile(JView jv, IView iv, double *__restrict__ part, uint32_t nact, uint32_t nact_pad,
             uint32_t tile_begin, uint32_t tile_end, uint32_t tiles_per_split, double e2)
{
    constexpr int TI = 32 * WI * IPT;   // i-particles per CTA
    constexpr int JW = TJ / WJ;         // j-particles of one tile handled by one warp
    static_assert(TJ % (2 * WJ) == 0, "tile must split into j pairs per warp");

    __shared__ __align__(128) double sj[NSTAGE][7][TJ];
    __shared__ __align__(8) uint64_t full[NSTAGE];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wi = warp % WI, wj = warp / WI;

    const uint32_t tb = tile_begin + blockIdx.y * tiles_per_split;
    const uint32_t te = min(tb + tiles_per_split, tile_end);
    const int ntile = te > tb ? (int)(te - tb) : 0;

    if (tid == 0)
    {
#pragma unroll
        for (int s = 0; s < NSTAGE; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncthreads();

    auto issue = [&](int t) {
        const int s = t % NSTAGE;
        const size_t off = (size_t)(tb + t) * TJ;
        mbar_expect_tx(&full[s], 7 * TJ * sizeof(double));
        bulk_g2s(&sj[s][0][0], jv.x + off, TJ * sizeof(double), &full[s]);
        bulk_g2s(&sj[s][1][0], jv.y + off, TJ * sizeof(double), &full[s]);
        bulk_g2s(&sj[s][2][0], jv.z + off, TJ * sizeof(double), &full[s]);
        bulk_g2s(&sj[s][3][0], jv.vx + off, TJ * sizeof(double), &full[s]);
        bulk_g2s(&sj[s][4][0], jv.vy + off, TJ * sizeof(double), &full[s]);
        bulk_g2s(&sj[s][5][0], jv.vz + off, TJ * sizeof(double), &full[s]);
        bulk_g2s(&sj[s][6][0], jv.m + off, TJ * sizeof(double), &full[s]);
    };
    if (tid == 0)
    {
        const int pre = ntile < NSTAGE ? ntile : NSTAGE;
        for (int t = 0; t < pre; ++t) issue(t);
    }

    // i-particles of this thread (indices past nact are clamped; their results are never stored)
    uint32_t k[IPT];
    double xi[IPT], yi[IPT], zi[IPT], vxi[IPT], vyi[IPT], vzi[IPT];
    Acc6 acc[IPT];
#pragma unroll
    for (int q = 0; q < IPT; ++q)
    {
        k[q] = blockIdx.x * TI + q * (32 * WI) + wi * 32 + lane;
        const uint32_t kc = k[q] < nact ? k[q] : nact - 1;
        xi[q] = iv.x[kc]; yi[q] = iv.y[kc]; zi[q] = iv.z[kc];
        vxi[q] = iv.vx[kc]; vyi[q] = iv.vy[kc]; vzi[q] = iv.vz[kc];
        acc[q] = Acc6{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    }

    for (int t = 0; t < ntile; ++t)
    {
        const int s = t % NSTAGE;
        mbar_wait(&full[s], (uint32_t)(t / NSTAGE) & 1u);
        const double *bx = &sj[s][0][wj * JW], *by = &sj[s][1][wj * JW], *bz = &sj[s][2][wj * JW];
        const double *bvx = &sj[s][3][wj * JW], *bvy = &sj[s][4][wj * JW], *bvz = &sj[s][5][wj * JW];
        const double *bm = &sj[s][6][wj * JW];
#pragma unroll 2
        for (int jj = 0; jj < JW; jj += 2)
        {
            const double2 X = *reinterpret_cast<const double2 *>(bx + jj);
            const double2 Y = *reinterpret_cast<const double2 *>(by + jj);
            const double2 Z = *reinterpret_cast<const double2 *>(bz + jj);
            const double2 VX = *reinterpret_cast<const double2 *>(bvx + jj);
            const double2 VY = *reinterpret_cast<const double2 *>(bvy + jj);
            const double2 VZ = *reinterpret_cast<const double2 *>(bvz + jj);
            const double2 M = *reinterpret_cast<const double2 *>(bm + jj);
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pair<MASK>(acc[q], xi[q], yi[q], zi[q], vxi[q], vyi[q], vzi[q], X.x, Y.x, Z.x, VX.x,
                           VY.x, VZ.x, M.x, e2);
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pair<MASK>(acc[q], xi[q], yi[q], zi[q], vxi[q], vyi[q], vzi[q], X.y, Y.y, Z.y, VX.y,
                           VY.y, VZ.y, M.y, e2);
        }
        __syncthreads();   // every warp is done with stage s -> it may be refilled
        if (tid == 0 && t + NSTAGE < ntile) issue(t + NSTAGE);
    }

    double *out = part + (size_t)blockIdx.y * 6 * nact_pad;
    if (WJ == 1)
    {
#pragma unroll
        for (int q = 0; q < IPT; ++q)
            if (k[q] < nact)
            {
                out[0 * (size_t)nact_pad + k[q]] = acc[q].ax;
                out[1 * (size_t)nact_pad + k[q]] = acc[q].ay;
                out[2 * (size_t)nact_pad + k[q]] = acc[q].az;
                out[3 * (size_t)nact_pad + k[q]] = acc[q].jx;
                out[4 * (size_t)nact_pad + k[q]] = acc[q].jy;
                out[5 * (size_t)nact_pad + k[q]] = acc[q].jz;
            }
    }
    else
    {
        // fold the WJ j-warps: red[wj][q][c][wi*32+lane], aliased on the (now idle) tile ring
        double *red = &sj[0][0][0];
        static_assert(WJ * IPT * 6 * 32 * WI <= NSTAGE * 7 * TJ, "reduce buffer must fit the ring");
        __syncthreads();
#pragma unroll
        for (int q = 0; q < IPT; ++q)
        {
            double *r = red + ((size_t)(wj * IPT + q) * 6) * (32 * WI) + wi * 32 + lane;
            r[0 * 32 * WI] = acc[q].ax; r[1 * 32 * WI] = acc[q].ay; r[2 * 32 * WI] = acc[q].az;
            r[3 * 32 * WI] = acc[q].jx; r[4 * 32 * WI] = acc[q].jy; r[5 * 32 * WI] = acc[q].jz;
        }
        __syncthreads();
        if (wj == 0)
        {
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                if (k[q] < nact)
                {
#pragma unroll
                    for (int c = 0; c < 6; ++c)
                    {
                        double s = 0.0;
#pragma unroll
                        for (int w = 0; w < WJ; ++w)
                            s += red[((size_t)(w * IPT + q) * 6 + c) * (32 * WI) + wi * 32 + lane];
                        out[c * (size_t)nact_pad + k[q]] = s;
                    }
                }
        }
    }
}

// ------------------------------------------------------------------------------------------------------
template <int IL, int IPT, int WJ, bool MASK>
__global__ void __launch_bounds__(32 * WJ)
k_force_lane(JView jv, IView iv, double *__restrict__ part, uint32_t nact, uint32_t nact_pad,
             uint32_t jbeg, uint32_t jend, uint32_t j_per_split, double e2)
{
    constexpr int JL = 32 / IL;       // j-lanes per warp
    constexpr int TI = IL * IPT;      // i-particles per CTA
    constexpr int PJ = JL * WJ;       // j-particles in flight per CTA step
    constexpr int U = 4;              // independent j per lane per trip (memory-level parallelism)

    __shared__ double red[WJ][IPT][6][IL];

    const int tid = threadIdx.x, lane = tid & 31, wj = tid >> 5;
    const int il = lane % IL, jl = lane / IL;

    const uint32_t jb = jbeg + blockIdx.y * j_per_split;
    const uint32_t je = min(jb + j_per_split, jend);

    uint32_t k[IPT];
    double xi[IPT], yi[IPT], zi[IPT], vxi[IPT], vyi[IPT], vzi[IPT];
    Acc6 acc[IPT];
#pragma unroll
    for (int q = 0; q < IPT; ++q)
    {
        k[q] = blockIdx.x * TI + q * IL + il;
        const uint32_t kc = k[q] < nact ? k[q] : nact - 1;
        xi[q] = iv.x[kc]; yi[q] = iv.y[kc]; zi[q] = iv.z[kc];
        vxi[q] = iv.vx[kc]; vyi[q] = iv.vy[kc]; vzi[q] = iv.vz[kc];
        acc[q] = Acc6{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    }

    for (uint32_t base = jb; base < je; base += PJ * U)
    {
        double xj[U], yj[U], zj[U], vxj[U], vyj[U], vzj[U], mj[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const uint32_t j = base + (u * WJ + wj) * JL + jl;
            const bool ok = j < je;
            const uint32_t jc = ok ? j : jb;
            xj[u] = __ldg(jv.x + jc); yj[u] = __ldg(jv.y + jc); zj[u] = __ldg(jv.z + jc);
            vxj[u] = __ldg(jv.vx + jc); vyj[u] = __ldg(jv.vy + jc); vzj[u] = __ldg(jv.vz + jc);
            const double mm = __ldg(jv.m + jc);
            mj[u] = ok ? mm : 0.0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int q = 0; q < IPT; ++q)
                pair<MASK>(acc[q], xi[q], yi[q], zi[q], vxi[q], vyi[q], vzi[q], xj[u], yj[u], zj[u],
                           vxj[u], vyj[u], vzj[u], mj[u], e2);
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
    // fold the WJ warps in fixed order; IPT*6*IL outputs per CTA
    double *out = part + (size_t)blockIdx.y * 6 * nact_pad;
    for (int o = tid; o < IPT * 6 * IL; o += 32 * WJ)
    {
        const int l = o % IL, c = (o / IL) % 6, q = o / (6 * IL);
        const uint32_t kk = blockIdx.x * TI + q * IL + l;
        if (kk < nact)
        {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < WJ; ++w) s += red[w][q][c][l];
            out[c * (size_t)nact_pad + kk] = s;
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Variant table.  kind 0 = tile (j staged through shared memory), 1 = lane (j from global).
struct ForceVariant
{
    int kind;       // 0 tile, 1 lane
    int ti;         // i-particles per CTA
    int threads;    // CTA size
    int pj;         // lane kernels: j-particles per CTA trip (PJ*U); tile kernels: TJ
    const char *name;
};

enum : int
{
    FV_TILE_I256 = 0,   // IPT=2 WI=4 WJ=1
    FV_TILE_I128,       // IPT=1 WI=4 WJ=1
    FV_TILE_I64,        // IPT=2 WI=1 WJ=4
    FV_TILE_I32,        // IPT=1 WI=1 WJ=4
    FV_LANE_I16,        // IL=16
    FV_LANE_I8,         // IL=8
    FV_LANE_I4,         // IL=4
    FV_LANE_I2,         // IL=2
    FV_LANE_I1,         // IL=1
    FV_LANE_I1x2,       // IL=1 IPT=2
    FV_LANE_I1x4,       // IL=1 IPT=4
    FV_TILE_I512,       // IPT=2 WI=8 WJ=1
    FV_COUNT
};

static const ForceVariant g_variants[FV_COUNT] = {
    {0, 256, 128, TJ, "tile_i256"}, {0, 128, 128, TJ, "tile_i128"}, {0, 64, 128, TJ, "tile_i64"},
    {0, 32, 128, TJ, "tile_i32"},   {1, 16, 128, 2 * 4 * 4, "lane_i16"}, {1, 8, 128, 4 * 4 * 4, "lane_i8"},
    {1, 4, 128, 8 * 4 * 4, "lane_i4"}, {1, 2, 128, 16 * 4 * 4, "lane_i2"}, {1, 1, 128, 32 * 4 * 4, "lane_i1"},
    {1, 2, 128, 32 * 4 * 4, "lane_i1x2"}, {1, 4, 128, 32 * 4 * 4, "lane_i1x4"}, {0, 512, 256, TJ, "tile_i512"},
};

template <bool MASK>
static cudaError_t launch_force(int variant, dim3 grid, cudaStream_t st, JView jv, IView iv,
                                double *part, uint32_t nact, uint32_t nact_pad, uint32_t jb,
                                uint32_t je, uint32_t per_split, double e2)
{
    // tile kernels take tile units (jb, je, per_split in tiles of TJ); lane kernels take particles
    switch (variant)
    {
    case FV_TILE_I256: k_force_tile<2, 4, 1, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_TILE_I128: k_force_tile<1, 4, 1, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_TILE_I64:  k_force_tile<2, 1, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_TILE_I32:  k_force_tile<1, 1, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_LANE_I16:  k_force_lane<16, 1, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_LANE_I8:   k_force_lane<8, 1, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_LANE_I4:   k_force_lane<4, 1, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_LANE_I2:   k_force_lane<2, 1, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_LANE_I1:   k_force_lane<1, 1, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_LANE_I1x2: k_force_lane<1, 2, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_LANE_I1x4: k_force_lane<1, 4, 4, MASK><<<grid, 128, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    case FV_TILE_I512: k_force_tile<2, 8, 1, MASK><<<grid, 256, 0, st>>>(jv, iv, part, nact, nact_pad, jb, je, per_split, e2); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace gdy
