// gdy_persist.cuh — several block steps per launch for SMALL active sets, without the host.
//
// More than half of all block steps move <= 8 particles.  Run one launch per hook they
// cost 6 dependent launches plus a host round trip each: ~45 us of latency for ~5 us of work.  This
// kernel is the reference loop body (reference: src/include/cpu/Hermite4CPU_integration.cpp:81-124:
// find_particles_to_move -> save_old_acc_jrk -> predicted_pos_vel -> update_acc_jrk ->
// correction_pos_vel -> next_integration_time) as ONE persistent cooperative kernel, one CTA per SM,
// that keeps stepping while the active set stays <= nact_max and hands back to the host loop
// (gdy_run) the moment a big block comes up.  Per block step:
//
//   S0  every CTA gathers the (ordered) active list from the per-CTA lists of the previous step
//   P1  every CTA predicts ITS j-slice straight from the state arrays into shared memory (the
//       predictor fused into the force evaluation: nothing is written back, 112 B/particle read
//       once), evaluates the pair sums of all active i against it (i-lanes x j-lanes, 2 i per
//       thread, shuffle + shared-memory fold) and leaves one partial per CTA; on the side it takes
//       the minimum of t+dt over the NON-active particles of its scan slice            -> grid sync
//   P2  one warp per active particle sums the per-CTA partials in fixed order (multi-GPU: and the
//       per-GPU sums, exchanged through peer memory over NVLink, in rank order), then corrects the
//       particle: save-old, Hermite corrector, new timestep level                       -> grid sync
//   P3  next block time = min(non-active minima, new t+dt of the corrected ones); every CTA lists the
//       particles of its scan slice that are due at that time, in ascending order      -> grid sync
//
// The arithmetic of predictor, corrector, timestep criterion and active test is the SAME device code
// as the one-launch-per-hook kernels (predict_one / correct_one / is_active_at in gdy_step.cuh), so the
// schedule is bit-identical given identical forces; the pair sums are grouped differently (per CTA
// slice), which moves forces by a few 1e-16 relative like any other j-split.
#pragma once
#include "gdy_device.cuh"
#include "gdy_step.cuh"

namespace gdy {

constexpr int PS_T = 512;              // threads per CTA (one CTA per SM)
constexpr int PS_W = PS_T / 32;
constexpr int PS_JCH = 2048;           // j-particles staged per chunk: 7 x 2048 x 8 B = 112 KiB
constexpr int PS_ICH = 64;             // active particles per i-chunk (32 i-lanes x 2)
constexpr int PS_NACT_CAP = 512;       // hard cap of nact_max (partial buffer, shared lists)
constexpr int PS_GMAX = 512;           // max CTAs
constexpr int PS_MAX_WORLD = 8;

struct PersistCtl   // device + pinned host copy
{
    double atime;        // next block time (ITIME of the step that was NOT done)
    double last_itime;   // ITIME of the last completed block step
    uint64_t sum_nact;   // over the completed steps of this launch
    uint32_t nact_next;  // size of the active list left in State.move
    uint32_t steps;      // block steps completed by this launch
    uint32_t last_nact;
    uint32_t seq;        // written last
};

struct TraceRec { double itime; uint32_t nact; uint32_t pad; };

struct PersistArgs
{
    State s;
    uint32_t n;
    uint32_t jb, je;           // this shard's j-particles [jb, je)
    double e2, eta, t_stop;
    double entry_itime;        // ITIME of the first step (= ATIME on entry)
    uint32_t entry_nact;       // its active list is State.move[0 .. entry_nact)
    uint32_t max_steps, nact_max;
    uint32_t slice;            // scan-slice length per CTA (cap of a per-CTA list)
    uint32_t seq;
    double *part;              // [nact_max][6][gpad]
    uint32_t gpad;
    double *cta_min;           // [G]
    uint32_t *cta_cnt;         // [G]
    uint32_t *cta_list;        // [G][slice]
    double *act_next;          // [nact_max]  t+dt of the corrected particles
    uint32_t *bar;             // grid barrier counter (zeroed before the launch)
    PersistCtl *ctl_dev, *ctl_host;
    TraceRec *trace;           // [max_steps]
    // multi-GPU exchange through peer memory (world == 1: unused)
    int rank, world;
    uint32_t xseq0;            // exchange sequence number of the first step of this launch
    double *xslot[PS_MAX_WORLD];      // rank r's buffer: [2][world][nact_max][6]
    uint32_t *xflag[PS_MAX_WORLD];    // rank r's flags : [2][world][nact_max]
};

// ---- grid barrier (all CTAs co-resident: cooperative launch) ---------------------------------------------
__device__ __forceinline__ void grid_sync(uint32_t *bar, uint32_t &target, uint32_t G)
{
    __syncthreads();
    if (threadIdx.x == 0)
    {
        target += G;
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar) : "memory");
        uint32_t v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
        } while (v < target);
        __threadfence();
    }
    __syncthreads();
}

// state that other SMs rewrite during the kernel's lifetime is read past L1
__device__ __forceinline__ double ldc(const double *p) { return __ldcg(p); }
__device__ __forceinline__ uint32_t ldc(const uint32_t *p) { return __ldcg(p); }

__device__ __forceinline__ Pred6 predict_one_cg(const State &s, uint32_t i, double itime)
{
    return predict_vals(itime, ldc(s.t + i), ldc(s.x + i), ldc(s.y + i), ldc(s.z + i), ldc(s.vx + i),
                        ldc(s.vy + i), ldc(s.vz + i), ldc(s.f[0] + i), ldc(s.f[1] + i), ldc(s.f[2] + i),
                        ldc(s.f[3] + i), ldc(s.f[4] + i), ldc(s.f[5] + i));
}

// pair sums of one i-chunk against the staged j-chunk.  Thread = (il, jl): i-lane il of IL, j-lane jl of
// PS_T/IL; two i per thread (k = kbase + il and kbase + IL + il).
template <int IL, bool MASK>
__device__ __forceinline__ void persist_pairs(const double *__restrict__ sj, int cnt, const double *si,
                                              int kloc0, double e2, Acc6 &a0, Acc6 &a1)
{
    constexpr int JL = PS_T / IL;
    const int il = threadIdx.x % IL, jl = threadIdx.x / IL;
    const double x0 = si[0 * PS_ICH + kloc0 + il], y0 = si[1 * PS_ICH + kloc0 + il], z0 = si[2 * PS_ICH + kloc0 + il];
    const double u0 = si[3 * PS_ICH + kloc0 + il], v0 = si[4 * PS_ICH + kloc0 + il], w0 = si[5 * PS_ICH + kloc0 + il];
    const double x1 = si[0 * PS_ICH + kloc0 + IL + il], y1 = si[1 * PS_ICH + kloc0 + IL + il], z1 = si[2 * PS_ICH + kloc0 + IL + il];
    const double u1 = si[3 * PS_ICH + kloc0 + IL + il], v1 = si[4 * PS_ICH + kloc0 + IL + il], w1 = si[5 * PS_ICH + kloc0 + IL + il];
#pragma unroll 2
    for (int j = jl; j < cnt; j += JL)
    {
        const double xj = sj[j], yj = sj[PS_JCH + j], zj = sj[2 * PS_JCH + j];
        const double uj = sj[3 * PS_JCH + j], vj = sj[4 * PS_JCH + j], wj = sj[5 * PS_JCH + j];
        const double mj = sj[6 * PS_JCH + j];
        pair<MASK>(a0, x0, y0, z0, u0, v0, w0, xj, yj, zj, uj, vj, wj, mj, e2);
        pair<MASK>(a1, x1, y1, z1, u1, v1, w1, xj, yj, zj, uj, vj, wj, mj, e2);
    }
}

template <int IL>
__device__ __forceinline__ void persist_fold(Acc6 &a0, Acc6 &a1, double *red /* [PS_W][2][6][32] */)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int off = IL; off < 32; off <<= 1)
    {
        a0.ax += __shfl_xor_sync(0xffffffffu, a0.ax, off); a0.ay += __shfl_xor_sync(0xffffffffu, a0.ay, off);
        a0.az += __shfl_xor_sync(0xffffffffu, a0.az, off); a0.jx += __shfl_xor_sync(0xffffffffu, a0.jx, off);
        a0.jy += __shfl_xor_sync(0xffffffffu, a0.jy, off); a0.jz += __shfl_xor_sync(0xffffffffu, a0.jz, off);
        a1.ax += __shfl_xor_sync(0xffffffffu, a1.ax, off); a1.ay += __shfl_xor_sync(0xffffffffu, a1.ay, off);
        a1.az += __shfl_xor_sync(0xffffffffu, a1.az, off); a1.jx += __shfl_xor_sync(0xffffffffu, a1.jx, off);
        a1.jy += __shfl_xor_sync(0xffffffffu, a1.jy, off); a1.jz += __shfl_xor_sync(0xffffffffu, a1.jz, off);
    }
    // IL >= 32: every lane is its own i-lane (il = lane); IL < 32: lanes [0, IL) hold the warp's sums
    constexpr int ILW = IL < 32 ? IL : 32;
    if (lane < ILW)
    {
        double *r = red + (size_t)w * 2 * 6 * 32 + lane;
        r[0 * 32] = a0.ax; r[1 * 32] = a0.ay; r[2 * 32] = a0.az; r[3 * 32] = a0.jx; r[4 * 32] = a0.jy; r[5 * 32] = a0.jz;
        r[6 * 32] = a1.ax; r[7 * 32] = a1.ay; r[8 * 32] = a1.az; r[9 * 32] = a1.jx; r[10 * 32] = a1.jy; r[11 * 32] = a1.jz;
    }
}

template <bool MASK>
__device__ __forceinline__ void persist_pairs_dispatch(int IL, const double *sj, int cnt, const double *si,
                                                       double e2, Acc6 &a0, Acc6 &a1)
{
    switch (IL)
    {
    case 1: persist_pairs<1, MASK>(sj, cnt, si, 0, e2, a0, a1); break;
    case 2: persist_pairs<2, MASK>(sj, cnt, si, 0, e2, a0, a1); break;
    case 4: persist_pairs<4, MASK>(sj, cnt, si, 0, e2, a0, a1); break;
    case 8: persist_pairs<8, MASK>(sj, cnt, si, 0, e2, a0, a1); break;
    case 16: persist_pairs<16, MASK>(sj, cnt, si, 0, e2, a0, a1); break;
    default: persist_pairs<32, MASK>(sj, cnt, si, 0, e2, a0, a1); break;
    }
}

__device__ __forceinline__ void persist_fold_dispatch(int IL, Acc6 &a0, Acc6 &a1, double *red)
{
    switch (IL)
    {
    case 1: persist_fold<1>(a0, a1, red); break;
    case 2: persist_fold<2>(a0, a1, red); break;
    case 4: persist_fold<4>(a0, a1, red); break;
    case 8: persist_fold<8>(a0, a1, red); break;
    case 16: persist_fold<16>(a0, a1, red); break;
    default: persist_fold<32>(a0, a1, red); break;
    }
}

struct PersistSmem
{
    double sj[7 * PS_JCH];            // staged predicted j-chunk, rows x y z vx vy vz m
    double si[6 * PS_ICH];            // predicted i-chunk
    double red[PS_W * 2 * 6 * 32];    // cross-warp fold
    double dmin[PS_W];
    uint32_t ilist[PS_NACT_CAP];
    uint32_t cnt[PS_GMAX];            // per-CTA counts, then exclusive prefix
    uint32_t wcnt[PS_W];
    uint32_t total;
};

template <bool MASK>
__global__ void __launch_bounds__(PS_T, 1)
k_persist_steps(const PersistArgs a)
{
    extern __shared__ __align__(16) unsigned char ps_raw[];
    PersistSmem &sm = *reinterpret_cast<PersistSmem *>(ps_raw);
    const State &s = a.s;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t G = gridDim.x, cta = blockIdx.x;
    uint32_t bar_target = 0;

    // this CTA's slices: j-slice of the shard (force evaluation), scan slice of all N (scheduler)
    const uint32_t nj = a.je - a.jb;
    const uint32_t j0 = a.jb + (uint32_t)(((uint64_t)nj * cta) / G), j1 = a.jb + (uint32_t)(((uint64_t)nj * (cta + 1)) / G);
    const uint32_t s0 = min(cta * a.slice, a.n), s1 = min(s0 + a.slice, a.n);

    double itime = a.entry_itime;
    uint32_t nact = a.entry_nact;
    uint32_t steps = 0;
    uint64_t sum_nact = 0;
    double last_itime = itime;
    uint32_t last_nact = 0;
    bool first = true;

    for (;;)
    {
        // ---- S0: the ordered active list of this step ----------------------------------------------------
        if (first)
        {
            for (uint32_t k = tid; k < nact; k += PS_T) sm.ilist[k] = s.move[k];
            __syncthreads();
        }
        else
        {
            // exclusive scan of the per-CTA counts (G <= 512: one per thread)
            uint32_t c = tid < (int)G ? ldc(a.cta_cnt + tid) : 0u;
            uint32_t incl = c;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1)
            {
                const uint32_t v = __shfl_up_sync(0xffffffffu, incl, off);
                if (lane >= off) incl += v;
            }
            if (lane == 31) sm.wcnt[w] = incl;
            __syncthreads();
            uint32_t woff = 0, tot = 0;
#pragma unroll
            for (int ww = 0; ww < PS_W; ++ww) { if (ww < w) woff += sm.wcnt[ww]; tot += sm.wcnt[ww]; }
            sm.cnt[tid] = woff + incl - c;     // exclusive prefix
            if (tid == 0) sm.total = tot;
            __syncthreads();
            nact = sm.total;
            // my own segment of h_move, so that the host loop can take over at any step — unless this launch
            // ends on its step budget or t_stop: then h_move keeps the list of the last COMPLETED step, as the
            // reference leaves it (src/include/Hermite4.cpp:63-84), and the host re-runs the scheduler if it
            // goes on
            const bool stop_end = steps >= a.max_steps || !(last_itime < a.t_stop);
            if (!stop_end)
            {
                const uint32_t mycnt = (cta + 1 < G ? sm.cnt[cta + 1] : nact) - sm.cnt[cta];
                const uint32_t off = sm.cnt[cta];
                for (uint32_t p = tid; p < mycnt; p += PS_T) s.move[off + p] = ldc(a.cta_list + (size_t)cta * a.slice + p);
            }
            const bool stop = steps >= a.max_steps || !(last_itime < a.t_stop) || nact == 0 || nact > a.nact_max;
            if (stop) break;
            for (uint32_t c2 = w; c2 < G; c2 += PS_W)
            {
                const uint32_t off = sm.cnt[c2];
                const uint32_t cc = (c2 + 1 < G ? sm.cnt[c2 + 1] : nact) - off;
                for (uint32_t p = lane; p < cc; p += 32) sm.ilist[off + p] = ldc(a.cta_list + (size_t)c2 * a.slice + p);
            }
            __syncthreads();
        }
        first = false;

        // ---- P1a: minimum of t+dt over the non-active particles of my scan slice -------------------------
        {
            double m = __longlong_as_double(0x7ff0000000000000LL);
            for (uint32_t i = s0 + tid; i < s1; i += PS_T)
            {
                const double tt = add(ldc(s.t + i), ldc(s.dt + i));
                if (!(fabs(sub(itime, tt)) < DBL_EPS2)) m = fmin(m, tt);
            }
#pragma unroll
            for (int off = 16; off; off >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, off));
            if (lane == 0) sm.dmin[w] = m;
            __syncthreads();
            if (tid == 0)
            {
                double mm = sm.dmin[0];
#pragma unroll
                for (int ww = 1; ww < PS_W; ++ww) mm = fmin(mm, sm.dmin[ww]);
                a.cta_min[cta] = mm;
            }
        }

        // ---- P1b: pair sums, i-chunk by i-chunk -----------------------------------------------------------
        const uint32_t njc = (j1 - j0 + PS_JCH - 1) / PS_JCH;
        for (uint32_t kb = 0; kb < nact; kb += PS_ICH)
        {
            const int nk = (int)min((uint32_t)PS_ICH, nact - kb);
            int IL = 1;
            while (2 * IL < nk) IL <<= 1;
            __syncthreads();   // si / red / sj of the previous chunk are free
            if (tid < PS_ICH)
            {
                const int kk = tid < nk ? tid : nk - 1;   // lanes past the list compute a duplicate, never stored
                const Pred6 p = predict_one_cg(s, sm.ilist[kb + kk], itime);
                sm.si[0 * PS_ICH + tid] = p.x; sm.si[1 * PS_ICH + tid] = p.y; sm.si[2 * PS_ICH + tid] = p.z;
                sm.si[3 * PS_ICH + tid] = p.vx; sm.si[4 * PS_ICH + tid] = p.vy; sm.si[5 * PS_ICH + tid] = p.vz;
            }
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
                        const Pred6 p = predict_one_cg(s, j, itime);
                        sm.sj[q] = p.x; sm.sj[PS_JCH + q] = p.y; sm.sj[2 * PS_JCH + q] = p.z;
                        sm.sj[3 * PS_JCH + q] = p.vx; sm.sj[4 * PS_JCH + q] = p.vy; sm.sj[5 * PS_JCH + q] = p.vz;
                        sm.sj[6 * PS_JCH + q] = __ldg(s.m + j);
                    }
                }
                __syncthreads();
                persist_pairs_dispatch<MASK>(IL, sm.sj, cnt, sm.si, a.e2, a0, a1);
            }
            persist_fold_dispatch(IL, a0, a1, sm.red);
            __syncthreads();
            // outputs o = (q, c, il): fold the warps in fixed order -> part[k][c][cta]
            const int ILW = IL < 32 ? IL : 32;
            for (int o = tid; o < 2 * 6 * ILW; o += PS_T)
            {
                const int il = o % ILW, c = (o / ILW) % 6, q = o / (6 * ILW);
                const int kloc = q * IL + il;
                if (kloc < nk)
                {
                    double sum = 0.0;
                    for (int ww = 0; ww < PS_W; ++ww) sum += sm.red[(size_t)ww * 2 * 6 * 32 + (q * 6 + c) * 32 + il];
                    a.part[((size_t)(kb + kloc) * 6 + c) * a.gpad + cta] = sum;
                }
            }
        }
        grid_sync(a.bar, bar_target, G);

        // ---- P2: one warp per active particle: sum the partials, (exchange,) correct ------------------------
        for (uint32_t k = cta + G * w; k < nact; k += G * PS_W)
        {
            double f[6];
#pragma unroll
            for (int c = 0; c < 6; ++c)
            {
                const double *p = a.part + ((size_t)k * 6 + c) * a.gpad;
                double v = 0.0;
                for (uint32_t cc = lane; cc < G; cc += 32) v += ldc(p + cc);
#pragma unroll
                for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
                f[c] = v;
            }
            if (a.world > 1)
            {
                // one-shot exchange over NVLink: my sum goes into slot [parity][rank][k] of EVERY rank
                // (mine included), flagged with the step's sequence number; then the world sums are
                // added in rank order, so every rank gets the same bits.
                const uint32_t xs = a.xseq0 + steps + 1;   // never 0
                const uint32_t par = xs & 1u;
                const size_t slot = (((size_t)par * a.world + a.rank) * a.nact_max + k);
                for (int idx = lane; idx < 6 * a.world; idx += 32)
                {
                    const int r = idx / 6, c = idx % 6;
                    double v = f[0];
#pragma unroll
                    for (int cc = 1; cc < 6; ++cc) if (c == cc) v = f[cc];
                    asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(a.xslot[r] + slot * 6 + c), "d"(v) : "memory");
                }
                __threadfence_system();
                __syncwarp();
                if (lane < a.world)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.xflag[lane] + slot), "r"(xs) : "memory");
                if (lane < a.world)
                {
                    const uint32_t *fl = a.xflag[a.rank] + (((size_t)par * a.world + lane) * a.nact_max + k);
                    uint32_t v;
                    do {
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(fl) : "memory");
                    } while (v != xs);
                }
                __syncwarp();
#pragma unroll
                for (int c = 0; c < 6; ++c) f[c] = 0.0;
                for (int r = 0; r < a.world; ++r)
                {
                    const double *sl = a.xslot[a.rank] + (((size_t)par * a.world + r) * a.nact_max + k) * 6;
#pragma unroll
                    for (int c = 0; c < 6; ++c)
                    {
                        double v;
                        asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(sl + c) : "memory");
                        f[c] += v;
                    }
                }
            }
            if (lane == 0)
            {
                const uint32_t i = sm.ilist[k];
                const Pred6 p = predict_one_cg(s, i, itime);
                store_pred(s, i, p);   // h_p of the active particles stays what the reference leaves there
                const double ndt = correct_one(s, i, itime, a.eta, p, f);
                a.act_next[k] = add(itime, ndt);
            }
        }
        grid_sync(a.bar, bar_target, G);

        // ---- P3: next block time and the per-CTA lists of the particles due then ----------------------------
        sum_nact += nact;
        last_itime = itime;
        last_nact = nact;
        if (cta == 0 && tid == 0 && steps < a.max_steps) a.trace[steps] = TraceRec{itime, nact, 0};
        ++steps;
        {
            double m = __longlong_as_double(0x7ff0000000000000LL);
            if (tid < (int)G) m = ldc(a.cta_min + tid);
            for (uint32_t k = tid; k < nact; k += PS_T) m = fmin(m, ldc(a.act_next + k));
#pragma unroll
            for (int off = 16; off; off >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, off));
            if (lane == 0) sm.dmin[w] = m;
            __syncthreads();
            m = sm.dmin[0];
#pragma unroll
            for (int ww = 1; ww < PS_W; ++ww) m = fmin(m, sm.dmin[ww]);
            itime = m;   // ATIME: ITIME of the next step
        }
        {
            // ordered compaction of my scan slice: warp ww owns a contiguous run; count, then write
            const uint32_t len = s1 - s0;
            const uint32_t per = ((len + PS_W - 1) / PS_W + 31) / 32 * 32;
            const uint32_t b = min(s0 + w * per, s1), e = min(b + per, s1);
            uint32_t cntw = 0;
            for (uint32_t base = b; base < e; base += 32)
            {
                const uint32_t i = base + lane;
                const bool on = i < e && fabs(sub(itime, add(ldc(s.t + i), ldc(s.dt + i)))) < DBL_EPS2;
                cntw += __popc(__ballot_sync(0xffffffffu, on));
            }
            if (lane == 0) sm.wcnt[w] = cntw;
            __syncthreads();
            uint32_t off = 0, tot = 0;
#pragma unroll
            for (int ww = 0; ww < PS_W; ++ww) { if (ww < w) off += sm.wcnt[ww]; tot += sm.wcnt[ww]; }
            uint32_t *list = a.cta_list + (size_t)cta * a.slice;
            for (uint32_t base = b; base < e; base += 32)
            {
                const uint32_t i = base + lane;
                const bool on = i < e && fabs(sub(itime, add(ldc(s.t + i), ldc(s.dt + i)))) < DBL_EPS2;
                const uint32_t bal = __ballot_sync(0xffffffffu, on);
                if (on) list[off + __popc(bal & ((1u << lane) - 1u))] = i;
                off += __popc(bal);
            }
            if (tid == 0) a.cta_cnt[cta] = tot;
        }
        grid_sync(a.bar, bar_target, G);
    }

    // ---- hand back to the host loop --------------------------------------------------------------------------
    if (cta == 0 && tid == 0)
    {
        PersistCtl c;
        c.atime = itime; c.last_itime = last_itime; c.sum_nact = sum_nact; c.nact_next = nact;
        c.steps = steps; c.last_nact = last_nact; c.seq = a.seq;
        *a.ctl_dev = c;
        volatile PersistCtl *h = a.ctl_host;
        h->atime = c.atime; h->last_itime = c.last_itime; h->sum_nact = c.sum_nact; h->nact_next = c.nact_next;
        h->steps = c.steps; h->last_nact = c.last_nact;
        __threadfence_system();
        h->seq = c.seq;
    }
}

}  // namespace gdy
