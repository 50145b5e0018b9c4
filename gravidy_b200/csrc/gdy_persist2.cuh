// gdy_persist2.cuh — the persistent small-step kernel, second design: state-resident, no grid barrier.
//
// Same job as round 1's k_persist_steps (removed; its building blocks live on in gdy_persist.cuh): the reference loop
// body (reference: src/include/cpu/Hermite4CPU_integration.cpp:81-124) for runs of block steps with small active sets, one
// cooperative launch, one CTA per SM.  What the first design paid per step — two grid barriers and a pass over
// the raw state of all N particles in L2 — is gone:
//
//   * Every CTA OWNS a slice of the particles for the lifetime of the launch.  It keeps the slice's raw state
//     (t, r, v, a, j: 13 doubles) in shared memory as far as it fits (the rest of the slice is streamed from the
//     state arrays, which only this CTA reads or writes while the kernel runs), the predicted slice (7 rows,
//     mass row static) and t+dt of its scan slice.  A corrected particle is patched in place by its owner.
//   * CTAs talk through flagged slots only (value + the step's tag written in one store; the reader polls until
//     the tag is the step's): predicted active particles (owner -> everybody), per-CTA partial forces
//     (everybody -> owner), the new t+dt of the corrected particles and the slice minima / candidate lists of the
//     scheduler (everybody -> everybody).  No fence, no barrier: a slot is self-validating, and every buffer
//     is rewritten only after a dependency chain that passes through all its readers (see the phase comments).
//   * j-sharded over GPUs, a rank's CTAs hold the rank's shard only; the owner of an active particle sums the
//     per-CTA partials of its rank, exchanges the rank sums with the other ranks through peer memory over NVLink
//     (flagged slots again) and adds them in rank order, so all ranks correct with identical bits.  A particle
//     that is not in the rank's shard has a fixed owner CTA as well (a function of its index), so that its state
//     arrays are only ever touched by one CTA of the rank.
//
// Ownership makes every raw-state access CTA-local (ordered by __syncthreads), which is what lets the kernel
// run without a single memory fence.  Every wait is bounded by a %globaltimer watchdog (per wait: cleared on success): on expiry the CTA
// raises the abort word, every other poll sees it, all CTAs leave, and the host returns GDY_ERR_STATE
// (the reference's convention is to fail loudly: src/include/gpu/Hermite4GPU.cuh:43-50).
//
// Arithmetic of predictor, corrector, timestep criterion and active test: the same device functions as the
// one-launch-per-hook kernels (gdy_step.cuh), so the schedule is bit-identical given identical forces.
#pragma once
#include "gdy_persist.cuh"

namespace gdy {

constexpr int P2_OWN = 64;       // owned active particles a CTA corrects per round of P2
constexpr int P2_STASH = 16;
constexpr int P2_NPROF = 16;

struct Persist2Args
{
    State s;
    uint32_t n;
    uint32_t jb, je;           // this shard's j-particles [jb, je)
    double e2, eta, t_stop;
    double entry_itime;
    uint32_t entry_nact;
    uint32_t max_steps, nact_max;
    uint32_t seq;
    uint32_t tag0;             // step k of this launch carries tag0 + k + 1 (never 0, unique over the context's life)
    // flagged-slot buffers of this rank
    uint4 *ipub;               // [CAP][6]        predicted active particles
    uint4 *part;               // [CAP][6][gpad]  per-CTA partial forces
    uint4 *anext;              // [CAP]           t+dt of the corrected particles
    uint4 *cmin;               // [2][gpad]       slice minimum of t+dt over the particles not due now
    uint2 *ccnt;               // [2][gpad]       {how many attain it, tag}
    uint2 *clist;              // [2][G][lcap]    {which, tag}, ascending
    uint2 *ecnt;               // [gpad]          exit: {due particles of the scan slice, tag}
    uint32_t gpad, lcap;
    // shared-memory carve-up behind P2Fixed (bytes / elements, set by the host)
    uint32_t off_dyn;          // byte offset of sj
    uint32_t lf_pad;           // row stride of sj  (>= the longest force slice)
    uint32_t res_cap;          // resident raw particles per CTA
    uint32_t ls_cap;           // scan-slice capacity
    PersistCtl *ctl_dev, *ctl_host;
    TraceRec *trace;
    unsigned long long *prof;  // [P2_NPROF] (NULL = off)
    uint32_t prof_nact_max;    // only block steps with at most this many active particles are timed (slot 14 counts them)
    uint32_t *abort_word;      // 0 = fine; else (phase << 16 | cta) + 1 of the first poll that gave up
    unsigned long long timeout_ns;
    int rank, world;
    uint4 *xslot[PS_MAX_WORLD];   // rank r's exchange buffer: [2][PS_MAX_WORLD][PS_NACT_CAP][6] flagged slots
};

struct P2Fixed
{
    double si[6 * PS_ICH];
    union   // the cross-warp fold of P1 and the two merge lists of P3 are never live together
    {
        double red[PS_W * 2 * 6 * 32];
        struct { uint32_t lista[PS_NACT_CAP], listb[PS_NACT_CAP]; } l;
    } u;
    double fs[P2_OWN * 6];
    double dmin[PS_W];
    uint32_t ilist[PS_NACT_CAP];
    uint32_t cnt[PS_GMAX + 1];
    uint32_t wcnt[PS_W];
    uint16_t own[PS_NACT_CAP];    // k of the active particles this CTA owns
    uint32_t nown;
    // what the owner already has in hand in phase A, kept for the corrector (first P2_STASH owned particles)
    double odt[P2_STASH];         // the particle's current step
    Pred6 opred[P2_STASH];        // its prediction at ITIME
    double wmin[PS_W];            // phase B: per-warp minimum of t+dt and how many attain it; P3: per-warp minimum
    uint32_t wa[PS_W], wb[PS_W];  // P3: per-warp counts of the two next-list sources
    uint32_t big;                 // P3: some slice has more than two candidates
    unsigned long long prof[P2_NPROF];   // phase timers of CTA 0 (flushed to global memory once, at exit)
};

// the CTA whose slice [len*c/G, len*(c+1)/G) holds x  (x < len)
__host__ __device__ __forceinline__ uint32_t slice_owner(uint32_t x, uint32_t len, uint32_t G)
{
    return (uint32_t)((((uint64_t)x + 1) * G + len - 1) / len) - 1;
}
__host__ __device__ __forceinline__ uint32_t slice_begin(uint32_t len, uint32_t c, uint32_t G)
{
    return (uint32_t)(((uint64_t)len * c) / G);
}

// One wait's watchdog.  t0 = %globaltimer at the first slow look of the CURRENT wait (0 = not yet timed); every wait that
// succeeds calls watch_reset, so the timeout measures a single wait, never the age of the launch.
struct Watch { unsigned long long t0; uint32_t spins; };
__device__ __forceinline__ void watch_reset(Watch &w) { w.t0 = 0ull; w.spins = 0u; }

// true = give up (the watchdog expired here or somewhere else)
__device__ __forceinline__ bool watch_expired(Watch &w, const Persist2Args &a, uint32_t where)
{
    if ((++w.spins & 255u) != 0) return false;
    uint32_t ab;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(ab) : "l"(a.abort_word) : "memory");
    if (ab) return true;
    const unsigned long long t = gtimer();
    if (w.t0 == 0) { w.t0 = t; return false; }
    if (t - w.t0 > a.timeout_ns)
    {
        atomicCAS(a.abort_word, 0u, ((where << 16) | blockIdx.x) + 1u);
        return true;
    }
    return false;
}

__device__ __forceinline__ bool ll_wait(const uint4 *slot, uint32_t tag, double &v, Watch &w, const Persist2Args &a,
                                        uint32_t where)
{
    for (;;)
    {
        const uint4 r = ll_load(slot);
        if (r.y == tag && r.w == tag) { v = ll_value(r); watch_reset(w); return true; }
        if (watch_expired(w, a, where)) { v = 0.0; return false; }
    }
}
__device__ __forceinline__ bool ll_wait_sys(const uint4 *slot, uint32_t tag, double &v, Watch &w, const Persist2Args &a,
                                            uint32_t where)
{
    for (;;)
    {
        const uint4 r = ll_load_sys(slot);
        if (r.y == tag && r.w == tag) { v = ll_value(r); watch_reset(w); return true; }
        if (watch_expired(w, a, where)) { v = 0.0; return false; }
    }
}
// 8-byte flagged entries {value, tag}
__device__ __forceinline__ void ll2_store(uint2 *slot, uint32_t v, uint32_t tag)
{
    asm volatile("st.relaxed.gpu.global.v2.b32 [%0], {%1, %2};" ::"l"(slot), "r"(v), "r"(tag) : "memory");
}
__device__ __forceinline__ bool ll2_wait(const uint2 *slot, uint32_t tag, uint32_t &v, Watch &w, const Persist2Args &a,
                                         uint32_t where)
{
    for (;;)
    {
        uint2 r;
        asm volatile("ld.relaxed.gpu.global.v2.b32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(slot) : "memory");
        if (r.y == tag) { v = r.x; watch_reset(w); return true; }
        if (watch_expired(w, a, where)) { v = 0u; return false; }
    }
}

template <bool MASK>
__global__ void __launch_bounds__(PS_T, 1)
k_persist2(const Persist2Args a)
{
    extern __shared__ __align__(16) unsigned char p2_raw[];
    P2Fixed &sm = *reinterpret_cast<P2Fixed *>(p2_raw);
    double *sj = reinterpret_cast<double *>(p2_raw + a.off_dyn);   // [7][lf_pad] predicted slice, row 6 = mass
    double *rs = sj + (size_t)7 * a.lf_pad;                        // [13][res_cap] resident raw state
    double *tn = rs + (size_t)13 * a.res_cap;                      // [ls_cap] t+dt of the scan slice
    const State &s = a.s;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t G = gridDim.x, cta = blockIdx.x;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const int JS = (int)a.lf_pad;
    const uint32_t RC = a.res_cap;

    // the slices this CTA owns: force slice (of the shard), scan slice (of all N)
    const uint32_t nj = a.je - a.jb;
    const uint32_t fj0 = a.jb + slice_begin(nj, cta, G), fj1 = a.jb + slice_begin(nj, cta + 1, G);
    const uint32_t Lf = fj1 - fj0, R = min(Lf, RC);
    const uint32_t s0 = slice_begin(a.n, cta, G), s1 = slice_begin(a.n, cta + 1, G);
    const uint32_t Ls = s1 - s0;
    const uint32_t per = ((Ls + PS_W * 32 - 1) / (PS_W * 32)) * 32;   // scan sub-slice of a warp (multiple of 32)

    auto owner = [&](uint32_t i) -> uint32_t {
        return (i >= a.jb && i < a.je) ? slice_owner(i - a.jb, nj, G) : slice_owner(i, a.n, G);
    };
    auto get_raw = [&](uint32_t i) -> Raw13 {
        const uint32_t q = i - fj0;
        if (q < R)
        {
            Raw13 r;
            r.t = rs[q]; r.x = rs[RC + q]; r.y = rs[2 * RC + q]; r.z = rs[3 * RC + q];
            r.vx = rs[4 * RC + q]; r.vy = rs[5 * RC + q]; r.vz = rs[6 * RC + q];
            r.f0 = rs[7 * RC + q]; r.f1 = rs[8 * RC + q]; r.f2 = rs[9 * RC + q];
            r.f3 = rs[10 * RC + q]; r.f4 = rs[11 * RC + q]; r.f5 = rs[12 * RC + q];
            return r;
        }
        return load_raw(s, i);   // only this CTA writes particle i while the kernel runs (ownership)
    };

    double itime = a.entry_itime;
    uint32_t nact = a.entry_nact;
    uint32_t steps = 0;
    uint32_t tag = 0;
    // Lane 0 of the calling warp corrects active particle k (position `opos` of this CTA's owned list) with the summed
    // force f: save_old_acc_jrk and the force store of update_acc_jrk fused, Hermite corrector, new timestep level; the
    // resident copy is patched and the particle's new t+dt published.
    auto correct_owned = [&](uint32_t k, uint32_t opos, const double *f) {
        if (lane != 0) return;
        const uint32_t i = sm.ilist[k];
        const Raw13 ri = get_raw(i);
        const bool stashed = opos < (uint32_t)P2_STASH;
        CorrIn q;
        q.dt1 = stashed ? sm.odt[opos] : ldc(s.dt + i);
        q.oa[0] = ri.f0; q.oa[1] = ri.f1; q.oa[2] = ri.f2; q.oj[0] = ri.f3; q.oj[1] = ri.f4; q.oj[2] = ri.f5;
#pragma unroll
        for (int c = 0; c < 3; ++c) { q.fa[c] = f[c]; q.fj[c] = f[3 + c]; }
        s.old[0][i] = ri.f0; s.old[1][i] = ri.f1; s.old[2][i] = ri.f2;
        s.old[3][i] = ri.f3; s.old[4][i] = ri.f4; s.old[5][i] = ri.f5;
#pragma unroll
        for (int c = 0; c < 6; ++c) s.f[c][i] = f[c];
        q.p = stashed ? sm.opred[opos] : predict_raw(ri, itime);
        store_pred(s, i, q.p);   // h_p of the active particles stays what the reference leaves there
        const CorrOut co = correct_core(s, i, itime, a.eta, q);
        const uint32_t qq = i - fj0;
        if (qq < R)   // patch the resident copy
        {
            rs[qq] = itime; rs[RC + qq] = co.x; rs[2 * RC + qq] = co.y; rs[3 * RC + qq] = co.z;
            rs[4 * RC + qq] = co.vx; rs[5 * RC + qq] = co.vy; rs[6 * RC + qq] = co.vz;
#pragma unroll
            for (int c = 0; c < 6; ++c) rs[(7 + c) * RC + qq] = f[c];
        }
        // anext[k] is rewritten a step later, behind every CTA's P3 of this step (see A)
        ll_store(a.anext + k, add(itime, co.ndt), tag);
    };
    uint64_t sum_nact = 0;
    double last_itime = itime;
    uint32_t last_nact = 0;
    bool stop_end = false, bail = false;
    Watch wd{0ull, 0u};
    bool dead = false;
    const unsigned long long t_start = (cta == 0 && tid == 0) ? gtimer() : 0ull;
    const bool prof = a.prof != nullptr && cta == 0 && tid == 0;
    unsigned long long tp = prof ? gtimer() : 0ull;
    bool profs = prof;         // this step is timed
#define GDY_PROF2(slot) \
    do { if (profs) { const unsigned long long tn_ = gtimer(); sm.prof[slot] += tn_ - tp; tp = tn_; } } while (0)
    if (prof)
        for (int q = 0; q < P2_NPROF; ++q) sm.prof[q] = 0ull;

    // ---- load the slices ------------------------------------------------------------------------------------------
    for (uint32_t q = tid; q < Lf; q += PS_T)
    {
        const uint32_t j = fj0 + q;
        sj[6 * JS + q] = ldc(s.m + j);
        if (q < R)
        {
            const Raw13 r = load_raw(s, j);
            rs[q] = r.t; rs[RC + q] = r.x; rs[2 * RC + q] = r.y; rs[3 * RC + q] = r.z;
            rs[4 * RC + q] = r.vx; rs[5 * RC + q] = r.vy; rs[6 * RC + q] = r.vz;
            rs[7 * RC + q] = r.f0; rs[8 * RC + q] = r.f1; rs[9 * RC + q] = r.f2;
            rs[10 * RC + q] = r.f3; rs[11 * RC + q] = r.f4; rs[12 * RC + q] = r.f5;
        }
    }
    for (uint32_t q = tid; q < Ls; q += PS_T) tn[q] = add(ldc(s.t + s0 + q), ldc(s.dt + s0 + q));
    for (uint32_t k = tid; k < nact; k += PS_T) sm.ilist[k] = ldc(s.move + k);
    if (tid == 0) sm.nown = 0;
    __syncthreads();
    GDY_PROF2(0);

    for (;;)
    {
        tag = a.tag0 + steps + 1;
        const uint32_t par = tag & 1u;
        profs = prof && nact <= a.prof_nact_max;
        if (profs) { tp = gtimer(); sm.prof[14] += 1ull; }

        // ---- A: the owners publish the predicted active particles ------------------------------------------------
        // ipub[k] is rewritten a step later, after its owner has been through P3 of this step, which needs the
        // new t+dt of every corrected particle, i.e. every owner's sums, i.e. every CTA's partials, i.e. every
        // CTA's reads of ipub.
        for (uint32_t k = tid; k < nact; k += PS_T)
        {
            const uint32_t i = sm.ilist[k];
            if (owner(i) == cta)
            {
                const uint32_t pos = atomicAdd(&sm.nown, 1u);
                sm.own[pos] = (uint16_t)k;
                const Pred6 p = predict_raw(get_raw(i), itime);
                if (pos < (uint32_t)P2_STASH) { sm.opred[pos] = p; sm.odt[pos] = ldc(s.dt + i); }
                uint4 *o = a.ipub + (size_t)k * 6;
                ll_store(o + 0, p.x, tag); ll_store(o + 1, p.y, tag); ll_store(o + 2, p.z, tag);
                ll_store(o + 3, p.vx, tag); ll_store(o + 4, p.vy, tag); ll_store(o + 5, p.vz, tag);
            }
        }
        GDY_PROF2(1);

        // ---- B: predictor of the slice (into shared memory), scheduler scan of the scan slice -----------------------
        for (uint32_t q = tid; q < Lf; q += PS_T)
        {
            const Pred6 p = predict_raw(get_raw(fj0 + q), itime);
            sj[q] = p.x; sj[JS + q] = p.y; sj[2 * JS + q] = p.z;
            sj[3 * JS + q] = p.vx; sj[4 * JS + q] = p.vy; sj[5 * JS + q] = p.vz;
        }
        {
            // minimum of t+dt over the particles of the scan slice that are NOT due now, and who attains it:
            // find_particles_to_move + next_integration_time of the next step for this slice
            // (reference: src/include/Hermite4.cpp:63-102)
            const uint32_t b0 = min((uint32_t)w * per, Ls), b1 = min(b0 + per, Ls);
            double m = INF;
            for (uint32_t q = b0 + lane; q < b1; q += 32)
            {
                const double v = tn[q];
                if (!due_at(itime, v)) m = fmin(m, v);
            }
#pragma unroll
            for (int off = 16; off; off >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, off));
            // how many of this warp's sub-slice attain the WARP's minimum; one barrier, then every thread folds the 16
            // (minimum, count) pairs for itself
            uint32_t c = 0;
            if (m < INF)
                for (uint32_t q0 = b0; q0 < b1; q0 += 32)
                {
                    const uint32_t q = q0 + lane;
                    const bool on = q < b1 && tn[q] == m;   // m is not due now, so neither is an entry equal to it
                    c += __popc(__ballot_sync(0xffffffffu, on));
                }
            if (lane == 0) { sm.wmin[w] = m; sm.wcnt[w] = c; }
            __syncthreads();
            double mm = sm.wmin[0];
#pragma unroll
            for (int ww = 1; ww < PS_W; ++ww) mm = fmin(mm, sm.wmin[ww]);
            uint32_t woff = 0, tot = 0;
            if (mm < INF)
            {
#pragma unroll
                for (int ww = 0; ww < PS_W; ++ww)
                    if (sm.wmin[ww] == mm) { if (ww < w) woff += sm.wcnt[ww]; tot += sm.wcnt[ww]; }
            }
            if (!(m == mm)) c = 0;   // this warp's sub-slice does not attain the slice minimum
            // clist[par] is rewritten two steps later; its readers (P3 of this step, every CTA) are through by then:
            // a CTA reaches B of step+2 only after P3 of step+1, which needs every CTA's partials of step+1, which
            // they publish after their P3 of this step.
            uint2 *list = a.clist + ((size_t)par * G + cta) * a.lcap;
            if (c)
            {
                uint32_t pos = woff;
                for (uint32_t q0 = b0; q0 < b1; q0 += 32)
                {
                    const uint32_t q = q0 + lane;
                    const bool on = q < b1 && tn[q] == mm;
                    const uint32_t bal = __ballot_sync(0xffffffffu, on);
                    const uint32_t at = pos + __popc(bal & ((1u << lane) - 1u));
                    if (on && at < a.lcap) ll2_store(list + at, s0 + q, tag);
                    pos += __popc(bal);
                }
            }
            if (tid == 0)
            {
                ll_store(a.cmin + par * a.gpad + cta, mm, tag);
                ll2_store(a.ccnt + par * a.gpad + cta, tot, tag);
            }
        }
        GDY_PROF2(2);

        // ---- P1: pair sums of the active particles against the slice, i-chunk by i-chunk -----------------------------
        for (uint32_t kb = 0; kb < nact && !bail; kb += PS_ICH)
        {
            const int nk = (int)min((uint32_t)PS_ICH, nact - kb);
            int IL = 1;
            while (2 * IL < nk) IL <<= 1;
            const int nwa = min(PS_W, (int)((Lf * (uint32_t)IL + 31) / 32));
            if (kb) __syncthreads();   // si / red of the previous chunk are free
            if (tid < 6 * PS_ICH)
            {
                const int kk = tid / 6, c = tid % 6;
                double v;
                if (!ll_wait(a.ipub + (size_t)(kb + min(kk, nk - 1)) * 6 + c, tag, v, wd, a, 1)) dead = true;
                sm.si[c * PS_ICH + kk] = v;   // past the list: a duplicate, never stored
            }
            GDY_PROF2(3);
            if (__syncthreads_or(dead)) { bail = true; break; }
            Acc6 a0{0, 0, 0, 0, 0, 0}, a1{0, 0, 0, 0, 0, 0};
            if (w < nwa) persist_pairs_fold<MASK>(IL, sj, (int)Lf, sm.si, a.e2, a0, a1, sm.u.red, true, JS);
            __syncthreads();
            // outputs o = (q, c, il): fold the warps in fixed order -> part[k][c][cta].  part[k] is rewritten a step
            // later, after P3 of this step, which needs its owner's sum, i.e. the owner's reads of it.
            for (int o = tid; o < 2 * 6 * IL; o += PS_T)
            {
                const int il = o % IL, c = (o / IL) % 6, q = o / (6 * IL);
                const int kloc = q * IL + il;
                if (kloc < nk)
                {
                    double sum = 0.0;
#pragma unroll
                    for (int ww = 0; ww < PS_W; ++ww) if (ww < nwa) sum += sm.u.red[(size_t)ww * 2 * 6 * 32 + (q * 6 + c) * 32 + il];
                    ll_store(a.part + ((size_t)(kb + kloc) * 6 + c) * a.gpad + cta, sum, tag);
                }
            }
            GDY_PROF2(4);
        }
        if (bail) break;

        // ---- P2: the owners sum the partials, (exchange the rank sums,) correct ------------------------------------------
        __syncthreads();
        {
            const uint32_t nown = sm.nown;
            for (uint32_t o0 = 0; o0 < nown && !bail; o0 += P2_OWN)
            {
                const uint32_t no = min((uint32_t)P2_OWN, nown - o0);
                // P2a: one warp per (particle, component): the G per-CTA partials in a fixed order
                for (uint32_t idx = w; idx < no * 6; idx += PS_W)
                {
                    const uint32_t k = sm.own[o0 + idx / 6], c = idx % 6;
                    const uint4 *p = a.part + ((size_t)k * 6 + c) * a.gpad;
                    double v = 0.0;
                    if (G <= 160)
                    {
                        // all five loads of a lane in flight together, re-polled as a group
                        uint4 r[5];
                        bool ok, giveup = false;
                        do {
                            ok = true;
#pragma unroll
                            for (int q = 0; q < 5; ++q)
                            {
                                const uint32_t cc = q * 32 + lane;
                                if (cc < G) { r[q] = ll_load(p + cc); ok = ok && r[q].y == tag && r[q].w == tag; }
                            }
                            if (!ok && watch_expired(wd, a, 2)) { giveup = true; dead = true; }
                        } while (!__all_sync(0xffffffffu, ok || giveup));
                        if (!giveup) watch_reset(wd);
#pragma unroll
                        for (int q = 0; q < 5; ++q) if ((uint32_t)(q * 32 + lane) < G) v += ll_value(r[q]);
                    }
                    else
                        for (uint32_t cc = lane; cc < G; cc += 32)
                        {
                            double x;
                            if (!ll_wait(p + cc, tag, x, wd, a, 2)) dead = true;
                            v += x;
                        }
#pragma unroll
                    for (int off = 16; off; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
                    if (lane == 0) sm.fs[idx] = v;
                }
                if (__syncthreads_or(dead)) { bail = true; break; }
                // P2b: one warp per particle.  Sharded: the rank's sum goes to EVERY rank's exchange buffer now, the
                // receive comes in a second pass once ALL owned particles of ALL rounds are on their way — a warp that
                // waited for particle p before sending particle q could wait for a rank whose own p-owner is behind its
                // q (the owners and their order differ from rank to rank): a cycle.  Sends never block.
                for (uint32_t o = w; o < no; o += PS_W)
                {
                    const uint32_t k = sm.own[o0 + o];
                    double f[6];
#pragma unroll
                    for (int c = 0; c < 6; ++c) f[c] = sm.fs[o * 6 + c];
                    if (a.world > 1)
                    {
                        // One-shot all-to-all over NVLink (replaces ncclAllReduce and the reference's host-side sum
                        // over GPUs, src/include/gpu/Hermite4GPU.cu:512-530): flagged slots into buffer
                        // [parity][my rank][k] of EVERY rank (mine included).  Two parities: a rank is at most one step
                        // ahead of the slowest reader (it cannot finish step+1 without this rank's sums of step+1).
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
                                ll_store_sys(a.xslot[r] + (((size_t)par * PS_MAX_WORLD + a.rank) * PS_NACT_CAP + k) * 6 + c, v, tag);
                            }
                        }
                    }
                    else
                        correct_owned(k, o0 + o, f);
                }
                if (o0 + P2_OWN < nown) __syncthreads();   // fs is reused
            }
            if (a.world > 1 && !bail)
                for (uint32_t o = w; o < nown; o += PS_W)
                {
                    // receive: the world's sums of particle k, added in rank order (every rank gets the same bits)
                    const uint32_t k = sm.own[o];
                    const bool xprof = a.prof != nullptr && nact <= a.prof_nact_max;
                    const unsigned long long tx = xprof ? gtimer() : 0ull;
                    double xv[2] = {0.0, 0.0};
#pragma unroll
                    for (int u = 0; u < 2; ++u)
                    {
                        const int idx = u * 32 + lane;
                        if (idx < 6 * a.world)
                        {
                            const int r = idx / 6, c = idx % 6;
                            if (!ll_wait_sys(a.xslot[a.rank] + (((size_t)par * PS_MAX_WORLD + r) * PS_NACT_CAP + k) * 6 + c,
                                             tag, xv[u], wd, a, 3)) dead = true;
                        }
                    }
                    __syncwarp();
                    double f[6];
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
                    if (xprof && lane == 0) { atomicAdd(a.prof + 12, gtimer() - tx); atomicAdd(a.prof + 13, 1ull); }
                    correct_owned(k, o, f);
                }
        }
        if (__syncthreads_or(dead)) bail = true;
        if (bail) break;
        GDY_PROF2(5);

        // ---- P3: next block time and next active list, by every CTA for itself ---------------------------------
        sum_nact += nact;
        last_itime = itime;
        last_nact = nact;
        ++steps;
        if (nact <= (uint32_t)PS_T)
        {
            // The common case, one corrected particle per thread at most: the same result as the general path below with
            // five CTA barriers instead of ten (every hand-over has its own shared-memory words).
            double an1 = INF, cmin1 = INF;
            uint32_t ccnt1 = 0;
            const bool mine = (uint32_t)tid < nact;
            if (mine)
            {
                if (!ll_wait(a.anext + tid, tag, an1, wd, a, 4)) dead = true;
                const uint32_t qs = sm.ilist[tid] - s0;
                if (qs < Ls) tn[qs] = an1;   // the scan slice's copy of t+dt
            }
            if (tid < (int)G)
            {
                if (!ll_wait(a.cmin + par * a.gpad + tid, tag, cmin1, wd, a, 5)) dead = true;
                if (!ll2_wait(a.ccnt + par * a.gpad + tid, tag, ccnt1, wd, a, 5)) dead = true;
            }
            GDY_PROF2(6);
            double m = fmin(an1, cmin1);
            __syncwarp();
#pragma unroll
            for (int off = 16; off; off >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, off));
            if (lane == 0) sm.wmin[w] = m;
            if (__syncthreads_or(dead)) { bail = true; break; }                                   // barrier 1
            double tnext = sm.wmin[0];
#pragma unroll
            for (int ww = 1; ww < PS_W; ++ww) tnext = fmin(tnext, sm.wmin[ww]);
            // corrected particles that are due again at tnext (ascending, like ilist) / candidates of the slices whose
            // minimum is tnext (slice order)
            const bool da = mine && due_at(tnext, an1);
            const uint32_t ba = __ballot_sync(0xffffffffu, da);
            const uint32_t bc = (tid < (int)G && due_at(tnext, cmin1)) ? ccnt1 : 0u;
            uint32_t incl = bc;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1)
            {
                const uint32_t v = __shfl_up_sync(0xffffffffu, incl, off);
                if (lane >= off) incl += v;
            }
            if (lane == 31) sm.wb[w] = incl;
            if (lane == 0) sm.wa[w] = __popc(ba);
            if (tid == 0) sm.big = 0;
            __syncthreads();                                                                      // barrier 2
            uint32_t aoff = 0, na = 0, boff = 0, nb = 0;
#pragma unroll
            for (int ww = 0; ww < PS_W; ++ww)
            {
                if (ww < w) { aoff += sm.wa[ww]; boff += sm.wb[ww]; }
                na += sm.wa[ww]; nb += sm.wb[ww];
            }
            const uint32_t nnext = na + nb;
            stop_end = steps >= a.max_steps || !(last_itime < a.t_stop);
            if (cta == 0 && tid == 0) a.trace[steps - 1] = TraceRec{itime, nact, (uint32_t)(gtimer() - t_start)};
            const uint32_t mycand = boff + incl - bc;   // where this slice's candidates go in listb
            const uint32_t myold = mine ? sm.ilist[tid] : 0u;
            itime = tnext;
            nact = nnext;
            if (stop_end || nnext == 0 || nnext > a.nact_max) break;
            if (da) sm.u.l.lista[aoff + __popc(ba & ((1u << lane) - 1u))] = myold;
            if (tid < (int)G) sm.cnt[tid] = mycand;
            if (tid == 0) { sm.cnt[G] = nb; sm.nown = 0; }
            if (bc)
            {
                // most candidate lists hold one or two particles: the slice's own thread fetches those; longer lists
                // are finished by whole warps behind the barrier
                const uint2 *src = a.clist + ((size_t)par * G + tid) * a.lcap;
                uint32_t v0 = 0, v1 = 0;
                if (!ll2_wait(src, tag, v0, wd, a, 6)) dead = true;
                if (bc > 1 && !ll2_wait(src + 1, tag, v1, wd, a, 6)) dead = true;
                sm.u.l.listb[mycand] = v0;
                if (bc > 1) sm.u.l.listb[mycand + 1] = v1;
                if (bc > 2) sm.big = 1;
            }
            if (__syncthreads_or(dead)) { bail = true; break; }                                   // barrier 3
            if (sm.big)
            {
                for (uint32_t c2 = w; c2 < G; c2 += PS_W)
                {
                    const uint32_t off = sm.cnt[c2], cc = sm.cnt[c2 + 1] - off;
                    const uint2 *src = a.clist + ((size_t)par * G + c2) * a.lcap;
                    for (uint32_t p = 2 + lane; p < cc; p += 32)
                    {
                        uint32_t v;
                        if (!ll2_wait(src + p, tag, v, wd, a, 6)) dead = true;
                        sm.u.l.listb[off + p] = v;
                    }
                }
                if (__syncthreads_or(dead)) { bail = true; break; }
            }
            for (uint32_t e = tid; e < na; e += PS_T)
            {
                const uint32_t v = sm.u.l.lista[e];
                sm.ilist[e + lower_bound_u32(sm.u.l.listb, nb, v)] = v;
            }
            for (uint32_t e = tid; e < nb; e += PS_T)
            {
                const uint32_t v = sm.u.l.listb[e];
                sm.ilist[e + lower_bound_u32(sm.u.l.lista, na, v)] = v;
            }
            __syncthreads();                                                                      // barrier 4
            GDY_PROF2(7);
            continue;
        }
        double an[PS_NACT_CAP / PS_T];
#pragma unroll
        for (int q = 0; q < PS_NACT_CAP / PS_T; ++q)
        {
            const uint32_t k = q * PS_T + tid;
            an[q] = INF;
            if (k < nact)
            {
                if (!ll_wait(a.anext + k, tag, an[q], wd, a, 4)) dead = true;
                const uint32_t qs = sm.ilist[k] - s0;
                if (qs < Ls) tn[qs] = an[q];   // the scan slice's copy of t+dt
            }
        }
        double cmin = INF;
        uint32_t ccnt = 0;
        if (tid < (int)G)
        {
            if (!ll_wait(a.cmin + par * a.gpad + tid, tag, cmin, wd, a, 5)) dead = true;
            if (!ll2_wait(a.ccnt + par * a.gpad + tid, tag, ccnt, wd, a, 5)) dead = true;
        }
        GDY_PROF2(6);
        if (__syncthreads_or(dead)) { bail = true; break; }
        double m = cmin;
#pragma unroll
        for (int q = 0; q < PS_NACT_CAP / PS_T; ++q) m = fmin(m, an[q]);
        const double tnext = block_min_ps(m, sm);
        // corrected particles that are due again at tnext (ascending, like ilist)
        uint32_t na = 0;
#pragma unroll
        for (int q = 0; q < PS_NACT_CAP / PS_T; ++q)
            if ((uint32_t)q * PS_T < nact)
            {
                const uint32_t k = q * PS_T + tid;
                compact_round(k < nact && due_at(tnext, an[q]), k < nact ? sm.ilist[k] : 0u, sm.u.l.lista, na, sm);
            }
        // candidates of the slices whose minimum is tnext
        const uint32_t bc = (tid < (int)G && due_at(tnext, cmin)) ? ccnt : 0u;
        const uint32_t nb = block_excl_scan(bc, G, sm);
        const uint32_t nnext = na + nb;
        stop_end = steps >= a.max_steps || !(last_itime < a.t_stop);
        if (cta == 0 && tid == 0) a.trace[steps - 1] = TraceRec{itime, nact, (uint32_t)(gtimer() - t_start)};
        itime = tnext;
        nact = nnext;
        if (stop_end || nnext == 0 || nnext > a.nact_max) break;
        for (uint32_t c2 = w; c2 < G; c2 += PS_W)
        {
            const uint32_t off = sm.cnt[c2], cc = sm.cnt[c2 + 1] - off;
            const uint2 *src = a.clist + ((size_t)par * G + c2) * a.lcap;
            for (uint32_t p = lane; p < cc; p += 32)
            {
                uint32_t v;
                if (!ll2_wait(src + p, tag, v, wd, a, 6)) dead = true;
                sm.u.l.listb[off + p] = v;
            }
        }
        if (tid == 0) sm.nown = 0;
        if (__syncthreads_or(dead)) { bail = true; break; }
        for (uint32_t e = tid; e < na; e += PS_T)
        {
            const uint32_t v = sm.u.l.lista[e];
            sm.ilist[e + lower_bound_u32(sm.u.l.listb, nb, v)] = v;
        }
        for (uint32_t e = tid; e < nb; e += PS_T)
        {
            const uint32_t v = sm.u.l.listb[e];
            sm.ilist[e + lower_bound_u32(sm.u.l.lista, na, v)] = v;
        }
        __syncthreads();
        GDY_PROF2(7);
    }

    // ---- hand back to the host loop --------------------------------------------------------------------------
    // A big block (or an empty one) comes next: leave its ordered list in State.move, scan slice by scan slice
    // (t+dt of the slice is in shared memory, patched).  When the launch ends on its step budget or on t_stop,
    // State.move keeps the list of the last COMPLETED step instead, as the reference leaves it
    // (src/include/Hermite4.cpp:63-84); the host re-runs the scheduler if it goes on.
    profs = prof;
    if (profs) tp = gtimer();
    if (!bail && !stop_end)
    {
        tag = a.tag0 + steps + 1;   // the host skips this tag
        const uint32_t b0 = min((uint32_t)w * per, Ls), b1 = min(b0 + per, Ls);
        uint32_t c = 0;
        for (uint32_t q0 = b0; q0 < b1; q0 += 32)
        {
            const uint32_t q = q0 + lane;
            c += __popc(__ballot_sync(0xffffffffu, q < b1 && due_at(itime, tn[q])));
        }
        __syncthreads();
        if (lane == 0) sm.wcnt[w] = c;
        __syncthreads();
        uint32_t woff = 0, tot = 0;
#pragma unroll
        for (int ww = 0; ww < PS_W; ++ww) { if (ww < w) woff += sm.wcnt[ww]; tot += sm.wcnt[ww]; }
        if (tid == 0) ll2_store(a.ecnt + cta, tot, tag);
        uint32_t ec = 0;
        if (tid < (int)G && !ll2_wait(a.ecnt + tid, tag, ec, wd, a, 7)) dead = true;
        block_excl_scan(ec, G, sm);
        if (__syncthreads_or(dead)) bail = true;
        else
        {
            uint32_t pos = sm.cnt[cta] + woff;
            for (uint32_t q0 = b0; q0 < b1; q0 += 32)
            {
                const uint32_t q = q0 + lane;
                const bool on = q < b1 && due_at(itime, tn[q]);
                const uint32_t bal = __ballot_sync(0xffffffffu, on);
                if (on) s.move[pos + __popc(bal & ((1u << lane) - 1u))] = s0 + q;
                pos += __popc(bal);
            }
        }
    }
    else if (!bail && cta == 0)
        for (uint32_t k = tid; k < last_nact; k += PS_T) s.move[k] = sm.ilist[k];   // the last completed step's list
    if (cta == 0 && tid == 0)
    {
        PersistCtl c;
        c.atime = itime; c.last_itime = last_itime; c.sum_nact = sum_nact; c.nact_next = nact;
        c.steps = steps; c.last_nact = last_nact; c.move_is_next = stop_end ? 0u : 1u; c.seq = a.seq; c.epoch_end = bail ? 1u : 0u;
        *a.ctl_dev = c;
        volatile PersistCtl *h = a.ctl_host;
        h->atime = c.atime; h->last_itime = c.last_itime; h->sum_nact = c.sum_nact; h->nact_next = c.nact_next;
        h->steps = c.steps; h->last_nact = c.last_nact; h->move_is_next = c.move_is_next; h->epoch_end = c.epoch_end;
        __threadfence_system();
        h->seq = c.seq;
    }
    GDY_PROF2(8);
    if (prof)
        for (int q = 0; q < P2_NPROF; ++q) if (q != 12 && q != 13) a.prof[q] += sm.prof[q];   // 12, 13: the owners' exchange waits (atomics)
#undef GDY_PROF2
}

}  // namespace gdy
