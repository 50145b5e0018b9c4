// gdy_persist.cuh — building blocks of the persistent small-step kernel (gdy_persist2.cuh): flagged 16-byte slots,
// the pair loop of an (i-lane x j-lane) CTA with its shuffle fold, ordered block compaction and scans.
//
// The first design of the kernel lived here (round 1: one CTA per SM, two grid barriers and one pass over the raw state
// of all N particles in L2 per block step, unbounded spin-waits).  It was replaced by the state-resident, barrier-free
// kernel of gdy_persist2.cuh (round 2: 16.7 instead of 23 us per small step at N = 256k on one GPU, 16 instead of 46-52 us
// on eight; profiles/r02_persist2_*.md) and removed.
#pragma once
#include "gdy_device.cuh"
#include "gdy_step.cuh"

namespace gdy {

constexpr int PS_T = 512;              // threads per CTA (one CTA per SM)
constexpr int PS_W = PS_T / 32;
constexpr int PS_JCH = 2048;           // j-particles staged per chunk: 7 x 2048 x 8 B = 112 KiB
constexpr int PS_ICH = 64;             // active particles per i-chunk (32 i-lanes x 2)
constexpr int PS_NACT_CAP = 2048;      // hard cap of nact_max (partial buffer, shared lists)
constexpr int PS_GMAX = 512;           // max CTAs (one flag / count per thread)
constexpr int PS_MAX_WORLD = 8;

struct PersistCtl   // device + pinned host copy
{
    double atime;        // next block time (ITIME of the step that was NOT done)
    double last_itime;   // ITIME of the last completed block step
    uint64_t sum_nact;   // over the completed steps of this launch
    uint32_t nact_next;  // size of the active set of the next step
    uint32_t steps;      // block steps completed by this launch
    uint32_t last_nact;
    uint32_t move_is_next;   // 1: State.move holds the NEXT step's list, 0: the last completed step's
    uint32_t seq;        // written last
    uint32_t epoch_end;  // grid-barrier epoch this launch ended on (the next launch continues from it)
};

struct TraceRec { double itime; uint32_t nact; uint32_t ns; };   // ns: end of the step, since the kernel started

// ---- flagged 16-byte slots (the "LL" protocol of collective libraries) --------------------------------------
// A double travels as {lo32, tag, hi32, tag}: each 8-byte half is written atomically WITH its tag, so the
// reader needs no fence and no barrier — it polls the slot until both tags carry the step's sequence number.
__device__ __forceinline__ void ll_store(uint4 *slot, double v, uint32_t tag)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    asm volatile("st.relaxed.gpu.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(slot), "r"((uint32_t)b), "r"(tag),
                 "r"((uint32_t)(b >> 32)), "r"(tag)
                 : "memory");
}
__device__ __forceinline__ uint4 ll_load(const uint4 *slot)
{
    uint4 r;
    asm volatile("ld.relaxed.gpu.global.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(slot) : "memory");
    return r;
}
// the same over NVLink (peer memory): system scope
__device__ __forceinline__ void ll_store_sys(uint4 *slot, double v, uint32_t tag)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    asm volatile("st.relaxed.sys.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(slot), "r"((uint32_t)b), "r"(tag),
                 "r"((uint32_t)(b >> 32)), "r"(tag)
                 : "memory");
}
__device__ __forceinline__ uint4 ll_load_sys(const uint4 *slot)
{
    uint4 r;
    asm volatile("ld.relaxed.sys.global.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(slot) : "memory");
    return r;
}
__device__ __forceinline__ double ll_value(const uint4 &r)
{
    return __longlong_as_double((long long)(((unsigned long long)r.z << 32) | (unsigned long long)r.x));
}

// state that other SMs rewrite during the kernel's lifetime is read past L1
__device__ __forceinline__ double ldc(const double *p) { return __ldcg(p); }
__device__ __forceinline__ uint32_t ldc(const uint32_t *p) { return __ldcg(p); }

struct Raw13 { double t, x, y, z, vx, vy, vz, f0, f1, f2, f3, f4, f5; };
__device__ __forceinline__ Raw13 load_raw(const State &s, uint32_t i)
{
    Raw13 r;
    r.t = ldc(s.t + i); r.x = ldc(s.x + i); r.y = ldc(s.y + i); r.z = ldc(s.z + i);
    r.vx = ldc(s.vx + i); r.vy = ldc(s.vy + i); r.vz = ldc(s.vz + i);
    r.f0 = ldc(s.f[0] + i); r.f1 = ldc(s.f[1] + i); r.f2 = ldc(s.f[2] + i);
    r.f3 = ldc(s.f[3] + i); r.f4 = ldc(s.f[4] + i); r.f5 = ldc(s.f[5] + i);
    return r;
}
__device__ __forceinline__ Pred6 predict_raw(const Raw13 &r, double itime)
{
    return predict_vals(itime, r.t, r.x, r.y, r.z, r.vx, r.vy, r.vz, r.f0, r.f1, r.f2, r.f3, r.f4, r.f5);
}

// the active test of find_particles_to_move (reference: src/include/Hermite4.cpp:75-77)
__device__ __forceinline__ bool due_at(double itime, double tt) { return fabs(sub(itime, tt)) < DBL_EPS2; }

// pair sums of one i-chunk against the staged j-chunk.  Thread = (il, jl): i-lane il of IL, j-lane jl of
// PS_T/IL; two i per thread (local k = il and IL + il).
template <int IL, bool MASK>
__device__ __forceinline__ void persist_pairs(const double *__restrict__ sj, int cnt, const double *si, double e2,
                                              Acc6 &a0, Acc6 &a1, const int JS = PS_JCH /* row stride of sj */)
{
    constexpr int JL = PS_T / IL;
    const int il = threadIdx.x % IL, jl = threadIdx.x / IL;
    const double x0 = si[0 * PS_ICH + il], y0 = si[1 * PS_ICH + il], z0 = si[2 * PS_ICH + il];
    const double u0 = si[3 * PS_ICH + il], v0 = si[4 * PS_ICH + il], w0 = si[5 * PS_ICH + il];
    const double x1 = si[0 * PS_ICH + IL + il], y1 = si[1 * PS_ICH + IL + il], z1 = si[2 * PS_ICH + IL + il];
    const double u1 = si[3 * PS_ICH + IL + il], v1 = si[4 * PS_ICH + IL + il], w1 = si[5 * PS_ICH + IL + il];
#pragma unroll 2
    for (int j = jl; j < cnt; j += JL)
    {
        const double xj = sj[j], yj = sj[JS + j], zj = sj[2 * JS + j];
        const double uj = sj[3 * JS + j], vj = sj[4 * JS + j], wj = sj[5 * JS + j];
        const double mj = sj[6 * JS + j];
        pair<MASK>(a0, x0, y0, z0, u0, v0, w0, xj, yj, zj, uj, vj, wj, mj, e2);
        pair<MASK>(a1, x1, y1, z1, u1, v1, w1, xj, yj, zj, uj, vj, wj, mj, e2);
    }
}

// Fold the 12 accumulators of a thread over the j-lanes of its warp (lane bits >= IL) by recursive halving:
// in every round a lane keeps half of its values and ships the other half to its partner, so 12 values
// cost 6+3+2+1+1 = 13 exchanges for IL = 1 instead of 12 x 5 with a butterfly per value.  Which value a
// lane ends up with is a fixed function of its lane bits; the order of the additions is fixed too.
template <int IL>
__device__ __forceinline__ void persist_fold(const Acc6 &a0, const Acc6 &a1, double *red /* [PS_W][12][32] */)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    double v[12] = {a0.ax, a0.ay, a0.az, a0.jx, a0.jy, a0.jz, a1.ax, a1.ay, a1.az, a1.jx, a1.jy, a1.jz};
    int id[12];
#pragma unroll
    for (int q = 0; q < 12; ++q) id[q] = q;
    int n = 12;
#pragma unroll
    for (int off = 16; off >= IL; off >>= 1)
    {
        const int half = (n + 1) / 2;
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int q = 0; q < 6; ++q)
            if (q < half)
            {
                const bool has_hi = q + half < n;
                const double lo = v[q], hi = has_hi ? v[q + half] : 0.0;
                const double recv = __shfl_xor_sync(0xffffffffu, up ? lo : hi, off);
                v[q] = (up ? hi : lo) + recv;
                id[q] = up ? (has_hi ? id[q + half] : -1) : id[q];
            }
        n = half;
    }
    double *r = red + (size_t)w * 12 * 32 + (lane & (IL - 1));
#pragma unroll
    for (int q = 0; q < 12; ++q)
        if (q < n && id[q] >= 0) r[id[q] * 32] = v[q];
}

template <bool MASK>
__device__ __forceinline__ void persist_pairs_fold(int IL, const double *sj, int cnt, const double *si, double e2,
                                                   Acc6 &a0, Acc6 &a1, double *red, bool fold, const int JS = PS_JCH)
{
    switch (IL)
    {
#define GDY_PS_CASE(n) \
    case n: persist_pairs<n, MASK>(sj, cnt, si, e2, a0, a1, JS); if (fold) persist_fold<n>(a0, a1, red); break;
        GDY_PS_CASE(1) GDY_PS_CASE(2) GDY_PS_CASE(4) GDY_PS_CASE(8) GDY_PS_CASE(16)
    default: persist_pairs<32, MASK>(sj, cnt, si, e2, a0, a1, JS); if (fold) persist_fold<32>(a0, a1, red); break;
#undef GDY_PS_CASE
    }
}

template <class SM>
__device__ __forceinline__ double block_min_ps(double m, SM &sm)
{
#pragma unroll
    for (int off = 16; off; off >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, off));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm.dmin[threadIdx.x >> 5] = m;
    __syncthreads();
    m = sm.dmin[0];
#pragma unroll
    for (int ww = 1; ww < PS_W; ++ww) m = fmin(m, sm.dmin[ww]);
    return m;
}

// one round of an ordered block compaction: flags `on` in thread order -> out[running + rank]
template <class SM>
__device__ __forceinline__ void compact_round(bool on, uint32_t value, uint32_t *out, uint32_t &running, SM &sm)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t bal = __ballot_sync(0xffffffffu, on);
    if (lane == 0) sm.wcnt[w] = __popc(bal);
    __syncthreads();
    uint32_t woff = 0, tot = 0;
#pragma unroll
    for (int ww = 0; ww < PS_W; ++ww) { if (ww < w) woff += sm.wcnt[ww]; tot += sm.wcnt[ww]; }
    if (on) out[running + woff + __popc(bal & ((1u << lane) - 1u))] = value;
    running += tot;
    __syncthreads();
}

// exclusive block scan of one value per thread t < G into sm.cnt[0 .. G), total in sm.cnt[G]  (G <= PS_T)
template <class SM>
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t c, uint32_t G, SM &sm)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = c;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1)
    {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += v;
    }
    __syncthreads();
    if (lane == 31) sm.wcnt[w] = incl;
    __syncthreads();
    uint32_t woff = 0, tot = 0;
#pragma unroll
    for (int ww = 0; ww < PS_W; ++ww) { if (ww < w) woff += sm.wcnt[ww]; tot += sm.wcnt[ww]; }
    if (threadIdx.x < G) sm.cnt[threadIdx.x] = woff + incl - c;
    if (threadIdx.x == 0) sm.cnt[G] = tot;
    __syncthreads();
    return tot;
}

__device__ __forceinline__ uint32_t lower_bound_u32(const uint32_t *a, uint32_t n, uint32_t v)
{
    uint32_t lo = 0, hi = n;
    while (lo < hi)
    {
        const uint32_t mid = (lo + hi) >> 1;
        if (a[mid] < v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ unsigned long long gtimer()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

}  // namespace gdy
