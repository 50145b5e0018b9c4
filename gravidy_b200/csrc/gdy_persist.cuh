// gdy_persist.cuh — several block steps per launch for SMALL active sets, without the host.
//
// More than half of all block steps move <= 8 particles.  Run one launch per hook they
// cost 6 dependent launches plus a host round trip each: ~50 us of latency for a few us of work.  This
// kernel is the reference loop body (reference: src/include/cpu/Hermite4CPU_integration.cpp:81-124:
// find_particles_to_move -> save_old_acc_jrk -> predicted_pos_vel -> update_acc_jrk ->
// correction_pos_vel -> next_integration_time) as ONE persistent cooperative kernel, one CTA per SM,
// that keeps stepping while the active set stays <= nact_max and hands back to the host loop
// (gdy_run) the moment a big block comes up.  A block step is two grid barriers:
//
//   P1  every CTA, for ITS slice of the particles: (a) the minimum of t+dt over the particles that are
//       NOT active now, and the ordered list of the particles that attain it — the slice's candidates
//       for the next block; (b) the predictor, straight from the state arrays into shared memory
//       (fused into the force evaluation: nothing is written back, 112 B/particle read once); (c) the
//       pair sums of all active i against the predicted slice (i-lanes x j-lanes, 2 i per thread,
//       shuffle + shared-memory fold) -> one partial per CTA                              -> barrier 1
//   P2  one warp per active particle sums the per-CTA partials in fixed order (multi-GPU: and the
//       per-GPU sums, exchanged through peer memory over NVLink, in rank order) and corrects the
//       particle: save-old, Hermite corrector, new timestep level, its new t+dt            -> barrier 2
//   P3  every CTA (redundantly, identical arithmetic): next block time = min(slice minima, new t+dt of
//       the corrected particles); next active list = merge of the corrected particles due then and of
//       the candidate lists of the slices whose minimum it is, ascending — find_particles_to_move and
//       next_integration_time without another pass over the particles.
//
// The arithmetic of predictor, corrector, timestep criterion and active test is the SAME device code
// as the one-launch-per-hook kernels (predict_vals / correct_core in gdy_step.cuh), so the schedule is
// bit-identical given identical forces; the pair sums are grouped differently (per CTA slice), which
// moves forces by a few 1e-16 relative like any other j-split.
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
constexpr int PS_DEFAULT_REPL_NACT = 16;   // sharded runs: blocks <= this are evaluated replicated, without exchange
constexpr int PS_RREG = 4;             // scan rounds whose t+dt stay in registers
constexpr int PS_NPROF = 16 + 2 * 64;   // 16 phase sums + two stamp rings (CTA 0, CTA G/2) of one step

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

struct PersistArgs
{
    State s;
    uint32_t n;
    uint32_t jb, je;           // this shard's j-particles [jb, je)
    double e2, eta, t_stop;
    double entry_itime;        // ITIME of the first step (= ATIME on entry)
    uint32_t entry_nact;       // its active list is State.move[0 .. entry_nact)
    uint32_t max_steps, nact_max;
    uint32_t slice;            // scan-slice length per CTA (cap of a per-CTA candidate list)
    uint32_t seq;
    uint4 *part;               // [nact_max][6][gpad] flagged slots, tag = sequence number of the step
    uint32_t gpad;
    double *cta_min;           // [2][G]      slice minima of t+dt over the non-active particles
    uint32_t *cta_cnt;         // [2][G]      how many particles of the slice attain it
    uint32_t *cta_list;        // [2][G][slice]  which, ascending
    uint32_t *exit_cnt;        // [G]
    double *act_next;          // [nact_max]  t+dt of the corrected particles
    uint32_t *flags;           // [G] grid-barrier epochs, all equal to epoch0 at entry
    uint32_t epoch0;
    PersistCtl *ctl_dev, *ctl_host;
    TraceRec *trace;           // [max_steps], pinned host memory (written through, never read here)
    unsigned long long *prof;  // [PS_NPROF] ns per phase, CTA 0 (NULL = off)
    // multi-GPU exchange through peer memory (world == 1: unused)
    int rank, world;
    uint32_t xseq0;            // exchange sequence number of the first step of this launch
    uint32_t xcount0;          // exchanges done before this launch (their parity picks the slot buffer)
    uint32_t repl_nact;        // world > 1: blocks of <= repl_nact active particles are evaluated against ALL N
                               // j-particles on every rank (replicated, no exchange): cheaper than the exchange
    uint4 *xslot[PS_MAX_WORLD];       // rank r's exchange buffer (peer memory): [2][PS_MAX_WORLD][PS_NACT_CAP][6] flagged slots
};

// ---- grid barrier (all CTAs co-resident: cooperative launch) ---------------------------------------------
// One epoch word per CTA: arrive = release-store of the epoch to my word (MEMBAR.ALL.GPU + store), wait =
// every thread t < G polls word t.  No atomics.  The waiting side issues no fence: everything this kernel
// reads from other CTAs is read past L1 (ld.global.cg / .relaxed.gpu = LDG.STRONG.GPU), after the poll's
// control dependency and the CTA barrier — an acquire fence would only add a CCTL.IVALL of the whole L1.
__device__ __forceinline__ void grid_barrier(uint32_t *flags, uint32_t epoch, uint32_t G)
{
    __syncthreads();
    if (threadIdx.x == 0)
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flags + blockIdx.x), "r"(epoch) : "memory");
    if (threadIdx.x < G)
    {
        uint32_t v;
        do {
            asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
        } while ((int32_t)(v - epoch) < 0);
    }
    __syncthreads();
}

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

struct PersistSmem
{
    double sj[7 * PS_JCH];            // staged predicted j-chunk, rows x y z vx vy vz m
    double si[6 * PS_ICH];            // predicted i-chunk
    double red[PS_W * 2 * 6 * 32];    // cross-warp fold
    double dmin[PS_W];
    double smin[PS_W];                // P1a, long slices: per-warp minimum of t+dt and how many attain it
    uint32_t scnt[PS_W];
    double fsum[2][6];                // P2: summed force of the two particles in flight
    uint32_t ilist[PS_NACT_CAP];      // active list of the current step
    uint32_t lista[PS_NACT_CAP];      // next: corrected particles that are due again
    uint32_t listb[PS_NACT_CAP];      // next: candidates of the slices
    uint32_t cnt[PS_GMAX + 1];        // exclusive prefix of per-CTA counts (+ total)
    uint32_t wcnt[PS_W];
};

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

// Warp-level scan of the particles [b0, b1): the minimum of t+dt over those NOT due at itime, and the ordered
// list of the particles that attain it (find_particles_to_move / next_integration_time of the next step,
// reference: src/include/Hermite4.cpp:63-102, for one slice).  The first PS_SCAN_RW rounds of 32 stay in
// registers; longer ranges are re-read for the list.
constexpr int PS_SCAN_RW = 8;
constexpr int PS_SCAN_SPLIT = 32;   // slices longer than this many rounds of 32 are scanned by all warps (measured:
                                    // N = 64k, 14 rounds: one warp hidden behind the staging is faster; N = 256k, 56: split is)
__device__ __forceinline__ void warp_scan_slice(const State &s, double itime, uint32_t b0, uint32_t b1, uint32_t *list,
                                                double &m_out, uint32_t &cnt_out)
{
    constexpr int RW = PS_SCAN_RW;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const int lane = threadIdx.x & 31;
    const uint32_t wr = (b1 - b0 + 31) / 32;
    double tt[RW];
#pragma unroll
    for (int r = 0; r < RW; ++r)
    {
        const uint32_t i = b0 + r * 32 + lane;
        tt[r] = i < b1 ? add(ldc(s.t + i), ldc(s.dt + i)) : INF;
    }
    double m = INF;
#pragma unroll
    for (int r = 0; r < RW; ++r) if (!due_at(itime, tt[r])) m = fmin(m, tt[r]);
    for (uint32_t r0 = RW; r0 < wr; r0 += RW)
    {
        double t2[RW];
#pragma unroll
        for (int r = 0; r < RW; ++r)
        {
            const uint32_t i = b0 + (r0 + r) * 32 + lane;
            t2[r] = i < b1 ? add(ldc(s.t + i), ldc(s.dt + i)) : INF;
        }
#pragma unroll
        for (int r = 0; r < RW; ++r) if (!due_at(itime, t2[r])) m = fmin(m, t2[r]);
    }
#pragma unroll
    for (int off = 16; off; off >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, off));
    uint32_t running = 0;
    if (m < INF)
    {
#pragma unroll
        for (int r = 0; r < RW; ++r)
            if ((uint32_t)r < wr)
            {
                const bool on = tt[r] == m && !due_at(itime, tt[r]);
                const uint32_t bal = __ballot_sync(0xffffffffu, on);
                if (on) list[running + __popc(bal & ((1u << lane) - 1u))] = b0 + r * 32 + lane;
                running += __popc(bal);
            }
        for (uint32_t r0 = RW; r0 < wr; r0 += RW)
        {
            double t2[RW];
#pragma unroll
            for (int r = 0; r < RW; ++r)
            {
                const uint32_t i = b0 + (r0 + r) * 32 + lane;
                t2[r] = i < b1 ? add(ldc(s.t + i), ldc(s.dt + i)) : INF;
            }
#pragma unroll
            for (int r = 0; r < RW; ++r)
            {
                const bool on = t2[r] == m && !due_at(itime, t2[r]);
                const uint32_t bal = __ballot_sync(0xffffffffu, on);
                if (on) list[running + __popc(bal & ((1u << lane) - 1u))] = b0 + (r0 + r) * 32 + lane;
                running += __popc(bal);
            }
        }
    }
    m_out = m;
    cnt_out = running;
}

template <bool MASK>
__global__ void __launch_bounds__(PS_T, 1)
k_persist_steps(const PersistArgs a)
{
    extern __shared__ __align__(16) unsigned char ps_raw[];
    PersistSmem &sm = *reinterpret_cast<PersistSmem *>(ps_raw);
    const State &s = a.s;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t G = gridDim.x, cta = blockIdx.x;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    uint32_t epoch = a.epoch0;

    // this CTA's slices: j-slice of the shard (force evaluation), scan slice of all N (scheduler)
    const uint32_t nj_sh = a.je - a.jb;
    const uint32_t j0_sh = a.jb + (uint32_t)(((uint64_t)nj_sh * cta) / G), j1_sh = a.jb + (uint32_t)(((uint64_t)nj_sh * (cta + 1)) / G);
    // ... and of ALL N, for the replicated tiny blocks of a sharded run (the same slices a one-GPU run uses)
    const uint32_t j0_all = (uint32_t)(((uint64_t)a.n * cta) / G), j1_all = (uint32_t)(((uint64_t)a.n * (cta + 1)) / G);
    const uint32_t s0 = min(cta * a.slice, a.n), s1 = min(s0 + a.slice, a.n);
    const uint32_t rounds = (s1 - s0 + PS_T - 1) / PS_T;
    uint32_t nx = a.xcount0;   // exchanges so far: identical on every rank

    double itime = a.entry_itime;
    uint32_t nact = a.entry_nact;
    uint32_t steps = 0;
    uint64_t sum_nact = 0;
    double last_itime = itime;
    uint32_t last_nact = 0;
    bool stop_end = false;
    const unsigned long long t_start = (cta == 0 && tid == 0) ? gtimer() : 0ull;
    const bool prof = a.prof != nullptr && cta == 0 && tid == 0;
    unsigned long long tp = prof ? gtimer() : 0ull;
#define GDY_PROF(slot) \
    do { if (prof) { const unsigned long long tn_ = gtimer(); a.prof[slot] += tn_ - tp; tp = tn_; } } while (0)
    const bool stamper = a.prof != nullptr && tid == 0 && (cta == 0 || cta == G / 2);
    unsigned long long *ring = a.prof + 16 + (cta == 0 ? 0 : 64);
    int nstamp = 0;
#define GDY_STAMP2(tag) \
    do { if (stamper && steps == 4 && nstamp < 32) { ring[2 * nstamp] = (unsigned long long)(tag); ring[2 * nstamp + 1] = gtimer(); ++nstamp; } } while (0)
#define GDY_STAMP(tag) \
    do { if (stamper && steps == 3 && nstamp < 32) { ring[2 * nstamp] = (unsigned long long)(tag); ring[2 * nstamp + 1] = gtimer(); ++nstamp; } } while (0)

    for (uint32_t k = tid; k < nact; k += PS_T) sm.ilist[k] = s.move[k];
    __syncthreads();

    for (;;)
    {
        const uint32_t par = steps & 1u;
        // uniform over the grid AND over the ranks (nact is replicated state)
        const bool repl = a.world > 1 && nact <= a.repl_nact;
        const uint32_t nj = repl ? a.n : nj_sh, j0 = repl ? j0_all : j0_sh, j1 = repl ? j1_all : j1_sh;
        const uint32_t njc = (j1 - j0 + PS_JCH - 1) / PS_JCH;
        const bool restage = (nj + G - 1) / G > (uint32_t)PS_JCH;
        GDY_STAMP(1);
        // ---- P1a: slice minimum of t+dt over the non-active particles and who attains it ----------------------
        // Short slices (<= 1024 particles): one warp does it alone, without CTA barriers, while the others already
        // stage the j-slice.  Longer slices (N > ~150k): every warp scans a contiguous sub-slice (one L2 latency
        // instead of a chain of them), candidates go to the sub-slice's own offset of the list, and one warp
        // closes the gaps of the sub-slices that attain the slice minimum.
        {
            const uint32_t wr = (s1 - s0 + 31) / 32;
            uint32_t *list = a.cta_list + ((size_t)par * G + cta) * a.slice;
            if (wr <= (uint32_t)PS_SCAN_SPLIT)
            {
                if (w == PS_W - 1)
                {
                    double m;
                    uint32_t running;
                    warp_scan_slice(s, itime, s0, s1, list, m, running);
                    if (lane == 0) { a.cta_min[par * G + cta] = m; a.cta_cnt[par * G + cta] = running; }
                }
            }
            else
            {
                const uint32_t per = ((wr + PS_W - 1) / PS_W) * 32;   // particles per warp
                const uint32_t b0 = min(s0 + (uint32_t)w * per, s1), b1 = min(b0 + per, s1);
                double m;
                uint32_t running;
                warp_scan_slice(s, itime, b0, b1, list + (b0 - s0), m, running);
                if (lane == 0) { sm.smin[w] = m; sm.scnt[w] = running; }
                __syncthreads();
                if (w == PS_W - 1)
                {
                    double mm = sm.smin[0];
#pragma unroll
                    for (int ww = 1; ww < PS_W; ++ww) mm = fmin(mm, sm.smin[ww]);
                    uint32_t tot = 0;
                    if (mm < INF)
                        for (int ww = 0; ww < PS_W; ++ww)
                            if (sm.smin[ww] == mm)
                            {
                                const uint32_t c = sm.scnt[ww];
                                const uint32_t so = min((uint32_t)ww * per, s1 - s0);
                                if (so != tot)   // move down: writes never pass the reads (tot <= so)
                                    for (uint32_t p0 = 0; p0 < c; p0 += 32)
                                    {
                                        const bool on = p0 + lane < c;
                                        const uint32_t v = on ? __ldcg(list + so + p0 + lane) : 0u;
                                        __syncwarp();
                                        if (on) list[tot + p0 + lane] = v;
                                        __syncwarp();
                                    }
                                tot += c;
                            }
                    if (lane == 0) { a.cta_min[par * G + cta] = mm; a.cta_cnt[par * G + cta] = tot; }
                }
            }
        }
        GDY_PROF(0);
        GDY_STAMP(4);

        // ---- P1b: pair sums, i-chunk by i-chunk -----------------------------------------------------------
        for (uint32_t kb = 0; kb < nact; kb += PS_ICH)
        {
            const int nk = (int)min((uint32_t)PS_ICH, nact - kb);
            int IL = 1;
            while (2 * IL < nk) IL <<= 1;
            if (kb) __syncthreads();   // si / red / sj of the previous chunk are free (kb = 0: P3 ended on a barrier)
            // warps whose first j-lane lies beyond the (longest) j-chunk have no pair to do: they skip the pair
            // loop and the fold, and the final sum runs over the others only
            const int nwa = min(PS_W, (int)((min((uint32_t)PS_JCH, j1 - j0) * (uint32_t)IL + 31) / 32));
            Raw13 ri{};
            if (tid < PS_ICH) ri = load_raw(s, sm.ilist[kb + (tid < nk ? tid : nk - 1)]);   // past the list: a duplicate, never stored
            GDY_STAMP(5);
            Acc6 a0{0, 0, 0, 0, 0, 0}, a1{0, 0, 0, 0, 0, 0};
            for (uint32_t jc = 0; jc < njc; ++jc)
            {
                const uint32_t c0 = j0 + jc * PS_JCH;
                const int cnt = (int)min((uint32_t)PS_JCH, j1 - c0);
                if (!(njc == 1 && kb > 0))
                {
                    if (jc > 0) __syncthreads();   // the previous j-chunk has been consumed
                    for (int q = tid; q < cnt; q += PS_T)
                    {
                        const uint32_t j = c0 + q;
                        const Raw13 rj = load_raw(s, j);
                        const double mj = __ldg(s.m + j);
                        const Pred6 p = predict_raw(rj, itime);
                        sm.sj[q] = p.x; sm.sj[PS_JCH + q] = p.y; sm.sj[2 * PS_JCH + q] = p.z;
                        sm.sj[3 * PS_JCH + q] = p.vx; sm.sj[4 * PS_JCH + q] = p.vy; sm.sj[5 * PS_JCH + q] = p.vz;
                        sm.sj[6 * PS_JCH + q] = mj;
                    }
                }
                GDY_STAMP(6);
                if (jc == 0 && tid < PS_ICH)
                {
                    const Pred6 p = predict_raw(ri, itime);
                    sm.si[0 * PS_ICH + tid] = p.x; sm.si[1 * PS_ICH + tid] = p.y; sm.si[2 * PS_ICH + tid] = p.z;
                    sm.si[3 * PS_ICH + tid] = p.vx; sm.si[4 * PS_ICH + tid] = p.vy; sm.si[5 * PS_ICH + tid] = p.vz;
                }
                GDY_PROF(8);
                GDY_STAMP(7);
                __syncthreads();
                GDY_PROF(9);
                GDY_STAMP(8);
                if (w < nwa) persist_pairs_fold<MASK>(IL, sm.sj, cnt, sm.si, a.e2, a0, a1, sm.red, jc + 1 == njc);
                GDY_PROF(10);
                GDY_STAMP(9);
            }
            __syncthreads();
            GDY_PROF(11);
            GDY_STAMP(10);
            // outputs o = (q, c, il): fold the warps in fixed order -> part[k][c][cta]
            for (int o = tid; o < 2 * 6 * IL; o += PS_T)
            {
                const int il = o % IL, c = (o / IL) % 6, q = o / (6 * IL);
                const int kloc = q * IL + il;
                if (kloc < nk)
                {
                    double sum = 0.0;
#pragma unroll
                    for (int ww = 0; ww < PS_W; ++ww) if (ww < nwa) sum += sm.red[(size_t)ww * 2 * 6 * 32 + (q * 6 + c) * 32 + il];
                    ll_store(a.part + ((size_t)(kb + kloc) * 6 + c) * a.gpad + cta, sum, a.xseq0 + steps + 1);
                }
            }
        }
        GDY_PROF(1);
        GDY_STAMP(11);
        // Slices longer than one j-chunk are re-staged from the state arrays for every i-chunk: then no
        // particle may be corrected before every CTA is through ALL its chunks (N > ~300k per GPU only).
        if (restage && nact > PS_ICH) grid_barrier(a.flags, ++epoch, G);

        // ---- P2: sum the partials as they arrive (no barrier: flagged slots), (exchange,) correct --------------
        // Two particles in flight per CTA: warps 0-5 / 6-11 poll one component each of particle kk / kk+G over
        // all G producers; warps 12 / 13 have the particle's state in flight meanwhile and do the correction.
        {
            const uint32_t tag = a.xseq0 + steps + 1;   // never 0
            for (uint32_t kk = cta; kk < nact; kk += 2 * G)
            {
                const int pslot = w < 12 ? w / 6 : w - 12;
                const uint32_t k = kk + pslot * G;
                const bool live = w < 14 && k < nact;
                Raw13 ri{};
                double dt1 = 0.0;
                uint32_t i = 0;
                if (live && w >= 12)
                {
                    i = sm.ilist[k];
                    if (lane == 0) { ri = load_raw(s, i); dt1 = ldc(s.dt + i); }
                }
                if (live && w < 12)
                {
                    const int c = w % 6;
                    const uint4 *p = a.part + ((size_t)k * 6 + c) * a.gpad;
                    double v = 0.0;
                    if (G <= 160)
                    {
                        uint4 r[5];
                        bool ok;
                        do {
                            ok = true;
#pragma unroll
                            for (int q = 0; q < 5; ++q)
                            {
                                const uint32_t cc = q * 32 + lane;
                                if (cc < G) { r[q] = ll_load(p + cc); ok = ok && r[q].y == tag && r[q].w == tag; }
                            }
                        } while (!__all_sync(0xffffffffu, ok));
#pragma unroll
                        for (int q = 0; q < 5; ++q) if ((uint32_t)(q * 32 + lane) < G) v += ll_value(r[q]);
                    }
                    else
                    {
                        for (uint32_t cc = lane; cc < G; cc += 32)
                        {
                            uint4 r;
                            do { r = ll_load(p + cc); } while (r.y != tag || r.w != tag);
                            v += ll_value(r);
                        }
                    }
#pragma unroll
                    for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
                    if (lane == 0) sm.fsum[pslot][c] = v;
                }
                __syncthreads();
                if (live && w >= 12)
                {
                    if (k == cta) GDY_STAMP(13);
                    double f[6];
#pragma unroll
                    for (int c = 0; c < 6; ++c) f[c] = sm.fsum[pslot][c];
                    if (a.world > 1 && !repl)
                    {
                        // One-shot all-to-all over NVLink, fused into the step (this replaces ncclAllReduce, and
                        // the reference's host-side sum over GPUs, src/include/gpu/Hermite4GPU.cu:512-530): my
                        // shard's sum goes as flagged slots into buffer [parity][my rank][k] of EVERY rank (mine
                        // included); then the world sums are added in rank order, so every rank gets the same
                        // bits.  Two parities: a rank can be at most one EXCHANGE ahead of the slowest reader (replicated
                        // steps in between do not synchronise the ranks, so the parity counts exchanges, not steps).
                        const uint32_t xp = nx & 1u;
                        double xv[2] = {0.0, 0.0};
#pragma unroll
                        for (int u = 0; u < 2; ++u)
                        {
                            const int idx = u * 32 + lane;
                            if (idx < 6 * a.world)
                            {
                                const int r = idx / 6, c = idx % 6;
                                double v = f[0];
#pragma unroll
                                for (int cc = 1; cc < 6; ++cc) if (c == cc) v = f[cc];
                                ll_store_sys(a.xslot[r] + (((size_t)xp * PS_MAX_WORLD + a.rank) * PS_NACT_CAP + k) * 6 + c, v, tag);
                            }
                        }
#pragma unroll
                        for (int u = 0; u < 2; ++u)
                        {
                            const int idx = u * 32 + lane;
                            if (idx < 6 * a.world)
                            {
                                const int r = idx / 6, c = idx % 6;
                                const uint4 *sl = a.xslot[a.rank] + (((size_t)xp * PS_MAX_WORLD + r) * PS_NACT_CAP + k) * 6 + c;
                                uint4 q;
                                do { q = ll_load_sys(sl); } while (q.y != tag || q.w != tag);
                                xv[u] = ll_value(q);
                            }
                        }
                        __syncwarp();
#pragma unroll
                        for (int c = 0; c < 6; ++c)
                        {
                            double sum = 0.0;
                            for (int r = 0; r < a.world; ++r)
                            {
                                const int idx = r * 6 + c;
                                const double v0 = __shfl_sync(0xffffffffu, xv[0], idx & 31);
                                const double v1 = __shfl_sync(0xffffffffu, xv[1], idx & 31);
                                sum += idx < 32 ? v0 : v1;
                            }
                            f[c] = sum;
                        }
                    }
                    if (lane == 0)
                    {
                        CorrIn q;
                        q.dt1 = dt1;
                        q.oa[0] = ri.f0; q.oa[1] = ri.f1; q.oa[2] = ri.f2; q.oj[0] = ri.f3; q.oj[1] = ri.f4; q.oj[2] = ri.f5;
#pragma unroll
                        for (int c = 0; c < 3; ++c) { q.fa[c] = f[c]; q.fj[c] = f[3 + c]; }
                        // save_old_acc_jrk and the force store of update_acc_jrk, fused
                        s.old[0][i] = ri.f0; s.old[1][i] = ri.f1; s.old[2][i] = ri.f2;
                        s.old[3][i] = ri.f3; s.old[4][i] = ri.f4; s.old[5][i] = ri.f5;
#pragma unroll
                        for (int c = 0; c < 6; ++c) s.f[c][i] = f[c];
                        q.p = predict_raw(ri, itime);
                        store_pred(s, i, q.p);   // h_p of the active particles stays what the reference leaves there
                        const double ndt = correct_core(s, i, itime, a.eta, q).ndt;
                        a.act_next[k] = add(itime, ndt);
                        if (k == cta) GDY_STAMP(14);
                    }
                }
                if (kk + 2 * G < nact) __syncthreads();   // fsum is reused
            }
        }
        if (cta == 0 && tid == 0) a.trace[steps] = TraceRec{itime, nact, (uint32_t)(gtimer() - t_start)};
        GDY_PROF(3);
        GDY_STAMP(15);
        grid_barrier(a.flags, ++epoch, G);
        GDY_PROF(4);
        GDY_STAMP(16);

        // ---- P3: next block time and next active list, by every CTA for itself ---------------------------------
        sum_nact += nact;
        last_itime = itime;
        last_nact = nact;
        ++steps;
        if (a.world > 1 && !repl) ++nx;
        double cmin = INF;
        uint32_t ccnt = 0;
        uint2 head = make_uint2(0u, 0u);   // first two candidates of slice `tid`: most lists end there
        if (tid < (int)G)
        {
            cmin = ldc(a.cta_min + par * G + tid); ccnt = ldc(a.cta_cnt + par * G + tid);
            head = __ldcg(reinterpret_cast<const uint2 *>(a.cta_list + ((size_t)par * G + tid) * a.slice));
        }
        double an[PS_NACT_CAP / PS_T];
#pragma unroll
        for (int q = 0; q < PS_NACT_CAP / PS_T; ++q)
        {
            const uint32_t k = q * PS_T + tid;
            an[q] = k < nact ? ldc(a.act_next + k) : INF;
        }
        double m = cmin;
#pragma unroll
        for (int q = 0; q < PS_NACT_CAP / PS_T; ++q) m = fmin(m, an[q]);
        const double tnext = block_min_ps(m, sm);
        GDY_STAMP2(17);
        // corrected particles that are due again at tnext (ascending, like ilist)
        uint32_t na = 0;
#pragma unroll
        for (int q = 0; q < PS_NACT_CAP / PS_T; ++q)
            if ((uint32_t)q * PS_T < nact)
            {
                const uint32_t k = q * PS_T + tid;
                compact_round(k < nact && due_at(tnext, an[q]), k < nact ? sm.ilist[k] : 0u, sm.lista, na, sm);
            }
        GDY_STAMP2(18);
        // candidates of the slices whose minimum is tnext
        const uint32_t bc = (tid < (int)G && due_at(tnext, cmin)) ? ccnt : 0u;
        const uint32_t nb = block_excl_scan(bc, G, sm);
        GDY_STAMP2(19);
        const uint32_t nnext = na + nb;
        stop_end = steps >= a.max_steps || !(last_itime < a.t_stop);
        itime = tnext;
        nact = nnext;
        if (stop_end || nnext == 0 || nnext > a.nact_max) break;
        if (bc >= 1) sm.listb[sm.cnt[tid]] = head.x;
        if (bc >= 2) sm.listb[sm.cnt[tid] + 1] = head.y;
        if (__syncthreads_or(bc > 2))
            for (uint32_t c2 = w; c2 < G; c2 += PS_W)
            {
                const uint32_t off = sm.cnt[c2], cc = sm.cnt[c2 + 1] - off;
                const uint32_t *src = a.cta_list + ((size_t)par * G + c2) * a.slice;
                if (cc > 2) for (uint32_t p = lane; p < cc; p += 32) sm.listb[off + p] = ldc(src + p);
            }
        __syncthreads();
        GDY_STAMP2(20);
        for (uint32_t e = tid; e < na; e += PS_T)
        {
            const uint32_t v = sm.lista[e];
            sm.ilist[e + lower_bound_u32(sm.listb, nb, v)] = v;
        }
        for (uint32_t e = tid; e < nb; e += PS_T)
        {
            const uint32_t v = sm.listb[e];
            sm.ilist[e + lower_bound_u32(sm.lista, na, v)] = v;
        }
        __syncthreads();
        GDY_PROF(5);
        GDY_STAMP2(21);
    }

    // ---- hand back to the host loop --------------------------------------------------------------------------
    // A big block (or an empty one) comes next: leave its ordered list in State.move, slice by slice.  When
    // the launch ends on its step budget or on t_stop, State.move keeps the list of the last COMPLETED step
    // instead, as the reference leaves it (src/include/Hermite4.cpp:63-84); the host re-runs the scheduler if
    // it goes on.
    if (!stop_end)
    {
        uint32_t c = 0;
        for (uint32_t r = 0; r < rounds; ++r)
        {
            const uint32_t i = s0 + r * PS_T + tid;
            const bool on = i < s1 && due_at(itime, add(ldc(s.t + i), ldc(s.dt + i)));
            c += __popc(__ballot_sync(0xffffffffu, on));
        }
        __syncthreads();
        if (lane == 0) sm.wcnt[w] = c;
        __syncthreads();
        if (tid == 0)
        {
            uint32_t tot = 0;
#pragma unroll
            for (int ww = 0; ww < PS_W; ++ww) tot += sm.wcnt[ww];
            a.exit_cnt[cta] = tot;
        }
        grid_barrier(a.flags, ++epoch, G);
        const uint32_t ec = tid < (int)G ? ldc(a.exit_cnt + tid) : 0u;
        block_excl_scan(ec, G, sm);
        uint32_t running = sm.cnt[cta];
        for (uint32_t r = 0; r < rounds; ++r)
        {
            const uint32_t i = s0 + r * PS_T + tid;
            const bool on = i < s1 && due_at(itime, add(ldc(s.t + i), ldc(s.dt + i)));
            compact_round(on, i, s.move, running, sm);
        }
    }
    else if (cta == 0)
        for (uint32_t k = tid; k < last_nact; k += PS_T) s.move[k] = sm.ilist[k];   // the last completed step's list
    if (cta == 0 && tid == 0)
    {
        PersistCtl c;
        c.atime = itime; c.last_itime = last_itime; c.sum_nact = sum_nact; c.nact_next = nact;
        c.steps = steps; c.last_nact = last_nact; c.move_is_next = stop_end ? 0u : 1u; c.seq = a.seq; c.epoch_end = epoch;
        *a.ctl_dev = c;
        volatile PersistCtl *h = a.ctl_host;
        h->atime = c.atime; h->last_itime = c.last_itime; h->sum_nact = c.sum_nact; h->nact_next = c.nact_next;
        h->steps = c.steps; h->last_nact = c.last_nact; h->move_is_next = c.move_is_next; h->epoch_end = c.epoch_end;
        __threadfence_system();
        h->seq = c.seq;
    }
    GDY_PROF(6);
#undef GDY_PROF
}

}  // namespace gdy
