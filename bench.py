#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on its config: Hermite FP64 GFLOP/s (60 flop/pair) at N = 256k.

A "step" is one pass of the hot path over one batch: the full-active acc+jerk evaluation
(init_acc_jrk's force sweep, N x N pairs including the self pair) of a seeded Plummer
model with N = 262144, j-sharded over the GPUs with the allreduce of the partial forces inside the
timed region.  Besides the contract's keys the JSON line carries

  roofline      the force kernel against the measured FP64 (DFMA) peak of this GPU
  cpu_baseline  the reference's own Hermite4CPU (oracle/_ref, else the C port) on the host cores, on a
                bounded sample of the same workload
  integration   a window of the N = 256k block-timestep integration (small active blocks): block steps/s,
                GFLOP/s and the extrapolated wall seconds per N-body time unit
  e2e           the same sweep through the C ABI with HOST buffers (upload r,v -> init_acc_jrk -> download
                forces), copies inside the timed region
  nbu           the second half of BASELINE.json's metric: the N = 256k block-timestep integration to t = 1 NBU,
                wall seconds per N-body time unit, block steps, sum of nact, cumulative energy error
  sweep_1m      the north-star target config: full-active sweep at N = 1 048 576, value and fraction of the FP64 peak
  checks        what the numbers rest on, verified in this very run (rc != 0 when one fails): a subset of the forces
                against the reference CPU code <= 1e-12; on several GPUs also: the forces after init_acc_jrk and the
                state after the integration window hash to the same value on every rank, and the window's schedule
                equals a one-GPU run of the same steps

`--impl reference` times the reference CPU implementation alone (rank 0 only).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_PAIR = 60.0
METRIC = "hermite_fp64_gflops_60flop_per_pair_N256k"
UNIT = "GFLOP/s"
NCU_DRAM_BYTES_PER_LAUNCH = 14886912 + 2654720   # profiles/r02_force_warp_ncu.md (capture r02u)


def workload_name(n):
    """config.workload, the same string in both arms (the driver compares them)."""
    return f"full-active acc+jerk sweep (init_acc_jrk force evaluation), Plummer N={n}, eps=1e-4, self pair counted"


def physical_cores():
    """Physical cores of this box (distinct (package, core) pairs of /proc/cpuinfo); None if it cannot be told."""
    try:
        seen, phys, core = set(), None, None
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("physical id"):
                phys = ln.split(":")[1].strip()
            elif ln.startswith("core id"):
                core = ln.split(":")[1].strip()
            elif not ln.strip():
                if phys is not None and core is not None:
                    seen.add((phys, core))
                phys = core = None
        return len(seen) or None
    except OSError:
        return None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", "--particles", dest="n", type=int, default=262144, help="particles (BASELINE config: 256k)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-integration", action="store_true")
    ap.add_argument("--integration-steps", type=int, default=400)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--cpu-loop-steps", type=int, default=100,
                    help="block steps of the N-body loop the reference CPU code is timed on (cpu_baseline.loop)")
    ap.add_argument("--no-nbu", action="store_true", help="skip the N=256k integration to t=1 (~100 s on one GPU)")
    ap.add_argument("--nbu-tend", type=float, default=1.0)
    ap.add_argument("--no-sweep-1m", action="store_true", help="skip the N=1M full-active sweep")
    ap.add_argument("--no-checks", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="only the timed sweep: no predictor/energy rooflines, e2e, integration, nbu, sweep_1m, checks")
    a = ap.parse_args()
    if a.no_extras:
        a.no_integration = a.no_nbu = a.no_sweep_1m = a.no_checks = a.no_cpu_baseline = True
    return a


# ---- CPU arm ----------------------------------------------------------------------------------------------
def cpu_force_rate(m, r, v, seconds, nthreads=None):
    """pairs/s of the reference's update_acc_jrk on a bounded subset of i-particles (all N j-particles)."""
    from oracle import oracle as O
    o = O.best_oracle(m, r, v)
    cores = nthreads or (os.cpu_count() or 1)
    o.set_threads(cores)
    n = len(m)
    rng = np.random.default_rng(12345)
    probe = np.sort(rng.choice(n, size=min(n, 4 * cores), replace=False)).astype(np.uint32)
    t = max(o.timed_update(probe), 1e-6)
    rate = len(probe) * n / t
    nact = int(min(n, max(4 * cores, rate * seconds / n)))
    nact -= nact % cores or 0
    nact = max(nact, cores)
    move = np.sort(rng.choice(n, size=nact, replace=False)).astype(np.uint32)
    t = o.timed_update(move)
    return nact * n / t, dict(kind=o.kind, cores=cores, nact=nact, seconds=t)


def run_reference(args, m, r, v):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = len(m)
    per_step = max(args.cpu_seconds / max(args.steps, 1), 2.0)
    for _ in range(args.warmup):
        cpu_force_rate(m, r, v, 0.5)
    rates, step_s, info = [], [], None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ts = time.perf_counter()
        rate, info = cpu_force_rate(m, r, v, per_step)
        step_s.append(time.perf_counter() - ts)
        rates.append(rate)
    wall = time.perf_counter() - t0
    rate = float(np.mean(rates))
    gflops = rate * FLOP_PER_PAIR / 1e9
    sample = (f"update_acc_jrk-style sweep of {info['nact']} of {n} i-particles x {n} j per step, "
              f"scaled to the full N^2")
    line = {
        "impl": "reference", "metric": METRIC, "value": gflops, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup,
        # a step of this arm IS the bounded sample (task contract): its measured duration, not a scaled one
        "ms_per_step": float(np.mean(step_s)) * 1e3,
        "ms_per_full_sweep_scaled": n * float(n) / rate * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(n), "n": n},
        "pairs_per_s": rate,
        "cpu_baseline": {"value": gflops, "unit": UNIT, "cores": info["cores"], "physical_cores": physical_cores(),
                         "kind": info["kind"], "sample": sample},
        "e2e": {"value": gflops, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": wall,
    }
    print(json.dumps(line), flush=True)


# ---- clocks ---------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """nvidia-smi-equivalent samples (NVML) of SM clock and throttle reasons during the timed region."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.power = index, [], set(), []
        self.stop_flag = threading.Event()
        self.max_mhz = None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for b, name in self.REASONS.items():
                    if bits & b:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "power_w_max": max(self.power) if self.power else None,
                "samples": len(self.samples)}


# ---- B200 arm ---------------------------------------------------------------------------------------------
def _sha(*arrays):
    import hashlib
    hh = hashlib.sha256()
    for a in arrays:
        hh.update(np.ascontiguousarray(a).tobytes())
    return hh.hexdigest()


def run_b200(args, m, r, v):
    import torch
    import torch.distributed as dist
    from gravidy_b200.hermite4 import Hermite4B200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; gravidy_b200 has no CPU path (use --impl reference "
                         "for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        tt = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    def gather(obj):
        if world == 1:
            return [obj]
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    def make_ctx(mm, rr, vv):
        """One j-shard per rank: NCCL communicator + peer memory for the fused NVLink exchanges."""
        uid = None
        if world > 1:
            t = torch.zeros(128, dtype=torch.uint8, device=dev)
            if rank == 0:
                t.copy_(torch.frombuffer(bytearray(Hermite4B200.nccl_unique_id()), dtype=torch.uint8))
            dist.broadcast(t, 0)
            uid = bytes(t.cpu().numpy().tobytes())
        hh = Hermite4B200(mm, rr, vv, softening=1e-4, eta=0.01, device=local, rank=rank, world=world, nccl_uid=uid)
        if world > 1:
            hh.peer_attach(gather(hh.peer_export()))
            dist.barrier()
        return hh

    checks, failed = {}, []

    def check(name, ok, detail=None):
        checks[name] = {"ok": bool(ok), **({"detail": detail} if detail is not None else {})}
        if not ok:
            failed.append(name)

    n = len(m)
    h = make_ctx(m, r, v)
    h.init_acc_jrk()
    h.sync()
    if not args.no_checks:
        f6_init = h.download_forces()
        hs = gather(_sha(f6_init))
        check("forces_after_init_acc_jrk_identical_on_all_ranks", len(set(hs)) == 1, hs[0][:16])
        if rank == 0:
            # 64 particles of the same sweep by the reference's own update_acc_jrk (oracle/_ref, else the C port)
            from oracle import oracle as O
            o = O.best_oracle(m, r, v)
            o.set_threads(os.cpu_count() or 1)
            sub = np.sort(np.random.default_rng(99).choice(n, size=64, replace=False)).astype(np.uint32)
            o.timed_update(sub)
            fr = o.get_state("f6")["f6"][sub]
            fg = f6_init[sub]
            ea = float((np.linalg.norm(fg[:, :3] - fr[:, :3], axis=1) / np.linalg.norm(fr[:, :3], axis=1)).max())
            ej = float((np.linalg.norm(fg[:, 3:] - fr[:, 3:], axis=1) / np.linalg.norm(fr[:, 3:], axis=1)).max())
            check("forces_vs_reference_cpu_64_particles_le_1e-12", ea <= 1e-12 and ej <= 1e-12,
                  {"acc_rel_err": ea, "jrk_rel_err": ej, "oracle": o.kind})
            o.close()
    peak_tf, peak_mhz = h.measure_fp64_peak()
    pairs = float(n) * float(n)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def one_step(hh=None):
        flush.zero_()
        barrier()
        return (hh or h).force_sweep_full(1)

    for _ in range(args.warmup):
        one_step()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = h.launch_count()
    barrier()
    w0 = time.perf_counter()
    ms = [one_step() for _ in range(args.steps)]
    barrier()
    wall_timed = time.perf_counter() - w0
    launches = h.launch_count() - l0
    sampler.stop_flag.set()
    sampler.join()
    total_ms = max_over_ranks(float(np.sum(ms)))
    ms_per_step = total_ms / args.steps
    rate = pairs / (ms_per_step * 1e-3)              # whole job: all ranks' pairs / max time
    gflops = rate * FLOP_PER_PAIR / 1e9

    # the dominant kernel alone (rank-local j-shard, no allreduce): same number at 1 GPU
    kern_tf = (pairs / world) * FLOP_PER_PAIR / (float(np.mean(ms)) * 1e-3) / 1e12

    hbm_peak, hbm_src = 6551.4, "fallback: the B200 copy bandwidth of MEASURED_PEAKS.json on this pool (file absent)"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            hbm_peak, hbm_src = float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (copy, read+write bytes)"
    except Exception:
        pass
    roofline_hbm = roofline_energy = None
    if not args.no_extras:
        # ---- the memory-bound kernel of the step: the predictor (120 B read + 48 B written per particle) ----
        pred_ms = []
        for _ in range(3 + 5):
            flush.zero_()
            torch.cuda.synchronize()
            pred_ms.append(h.predict_timed(0.0))
        pred_ms = float(np.mean(pred_ms[3:]))
        pred_gbs = 168.0 * n / (pred_ms * 1e-3) / 1e9
        roofline_hbm = {"bound": "hbm", "kernel": "k_predict (predicted_pos_vel, all N, L2 flushed before each launch)",
                        "achieved": pred_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": pred_gbs / hbm_peak,
                        "us_per_launch": pred_ms * 1e3, "algorithmic_bytes_per_particle": 168, "traffic": None,
                        "peak_source": hbm_src,
                        "note": "one event-bracketed launch of ~10 us: launch latency is a third of it"}
        # the same kernel where it is bandwidth-bound: N = 2M (352 MB per launch, nothing survives in the 126 MB L2),
        # launches back to back — what the kernel sustains once launch latency is amortised
        if rank == 0 and world == 1:
            nb = 2097152
            rng = np.random.default_rng(7)
            hb = Hermite4B200(np.full(nb, 1.0 / nb), rng.standard_normal((nb, 3)), rng.standard_normal((nb, 3)),
                              softening=1e-4, eta=0.01, device=local)
            for _ in range(3):
                hb.predict_timed(0.0)
            reps = 20
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            big_ms = float(np.mean([hb.predict_timed(0.125 * (k + 1)) for k in range(reps)]))
            big_wall = (time.perf_counter() - t0) / reps * 1e3
            hb.close()
            big_gbs = 168.0 * nb / (big_ms * 1e-3) / 1e9
            roofline_hbm["bandwidth_regime"] = {"n": nb, "us_per_launch": big_ms * 1e3, "achieved": big_gbs,
                                                "frac": big_gbs / hbm_peak, "wall_us_per_call": big_wall * 1e3,
                                                "note": "same k_predict at N = 2M, working set 352 MB > L2, "
                                                        "event-bracketed launches back to back"}
        # ---- the other two FP64-bound kernels of SURVEY 8(d): energy and the k-th neighbour search ----------
        if hasattr(h, "energy_timed"):
            roofline_energy = {}
            # k_energy is i-sharded over the ranks (each rank: its i-chunks x all j); the k-th-neighbour search serves the
            # host program's density centre on the rank that writes the log and is NOT sharded: every rank runs all N x N
            for nm, fn, instr, sharded in (("k_energy", h.energy_timed, 12.0, True),
                                           ("k_knn_distance", lambda: h.neighbour_timed(6), 9.0, False)):
                fn()
                tms = float(np.mean([fn() for _ in range(3)]))
                mine = pairs / world if sharded else pairs
                # the FP64 pipe retires 64 lanes per clock per SM: instr / pair against the DFMA peak's instruction rate
                ach = mine * instr / (tms * 1e-3) / 1e12               # T(FP64 instr)/s of this rank
                pk = peak_tf / 2.0                                      # the peak in DFMA instructions (2 flop each)
                roofline_energy[nm] = {"bound": "fp64", "ms_per_launch": tms, "fp64_instr_per_ordered_pair": instr,
                                       "achieved": ach, "peak": pk, "unit": "T FP64 instr/s", "frac": ach / pk,
                                       "pairs_per_s": mine / (tms * 1e-3),
                                       "i_slice": (f"rank's slice of {n // world} i-particles x all {n} j"
                                                   if world > 1 and sharded else "all N x N")}

    # ---- e2e: host buffers through the C ABI, copies inside the timed region --------------------------
    r4 = torch.empty((n, 4), dtype=torch.float64).pin_memory().numpy()
    v4 = torch.empty((n, 4), dtype=torch.float64).pin_memory().numpy()
    f6 = torch.empty((n, 6), dtype=torch.float64).pin_memory().numpy()
    r4[:, :3] = r; r4[:, 3] = m; v4[:, :3] = v; v4[:, 3] = 0.0
    e2e = None
    if not args.no_extras:
        # A sharded upload is collective: every rank is handed the same arrays, copies its own chunk over PCIe and the
        # chunks are all-gathered over NVLink; the step's result is read back where it is wanted — on rank 0.
        def e2e_step():
            h.upload_particles(r4, v4)
            h.init_acc_jrk()
            if rank == 0:
                h.download_forces(out=f6)
            else:
                h.sync()

        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        barrier()
        e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
        e2e = {"value": pairs / e2e_s * FLOP_PER_PAIR / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": int(r4.nbytes + v4.nbytes),   # whole job: each rank copies 1/world of it
               "d2h_bytes_per_step": int(f6.nbytes),
               "ms_per_step": e2e_s * 1e3,
               "call": "gdy_upload_particles + gdy_init_acc_jrk + gdy_download_forces (pinned host buffers)"
                       + (f"; each rank uploads its 1/{world} of the particles and the chunks are all-gathered over NVLink, "
                          "rank 0 downloads the forces" if world > 1 else "")}

    # ---- a window of the block-timestep integration (small active blocks) -----------------------------
    integ = None
    loop_head = start = None
    if not args.no_integration:
        at = h.init_dt(0.0)
        want_cpu_loop = rank == 0 and world == 1 and not args.no_cpu_baseline and args.cpu_loop_steps > 0
        start = h.get_state("f6", "t", "dt") if want_cpu_loop else None   # the loop's initial state, for the CPU arm
        k0 = max(args.cpu_loop_steps, 50)
        barrier()
        st, tn0, _, tu0 = h.run(at, 1e9, max_steps=k0, trace_cap=k0, trace_us=True)
        barrier()
        loop_head = {"atime0": at, "nact": tn0, "us": tu0}
        l1 = h.launch_count()
        x0 = h.exchange_stats()
        barrier()
        st, tn, _, tu = h.run(st.atime, 1e9, max_steps=args.integration_steps,
                              trace_cap=args.integration_steps, itime0=st.itime, trace_us=True)
        barrier()
        wall = max_over_ranks(st.wall_ms) * 1e-3
        x1 = h.exchange_stats()
        integ = {"workload": f"block-timestep integration, Plummer N={n}, block steps {k0}..{k0 + st.iterations}",
                 "block_steps": int(st.iterations), "sum_nact": int(st.nsteps),
                 "mean_nact": float(st.nsteps) / max(int(st.iterations), 1),
                 "median_nact": float(np.median(tn)) if len(tn) else None,
                 "pairs": int(st.interactions), "wall_s": wall,
                 "us_per_block_step": wall / max(int(st.iterations), 1) * 1e6,
                 "gflops": st.interactions / wall * FLOP_PER_PAIR / 1e9,
                 "launches": h.launch_count() - l1, "t_reached": st.itime,
                 "us_per_block_step_nact_le_32": float(np.median(tu[tn <= 32])) if (tn <= 32).any() else None,
                 "mean_us_per_block_step_nact_le_512": float(np.mean(tu[tn <= 512])) if (tn <= 512).any() else None,
                 "exchanges_peer_memory": x1[0] - x0[0], "exchanges_nccl": x1[1] - x0[1]}
        if not args.no_checks:
            stt = h.get_state("r4", "v4", "f6", "t", "dt")
            hs = gather(_sha(stt["r4"], stt["v4"], stt["f6"], stt["t"], stt["dt"]))
            check("state_after_integration_window_identical_on_all_ranks", len(set(hs)) == 1, hs[0][:16])
            if world > 1:
                # the same block steps on ONE GPU (rank 0, a second context): the schedule must be the same
                same = None
                if rank == 0:
                    h1 = Hermite4B200(m, r, v, softening=1e-4, eta=0.01, device=local)
                    h1.init_acc_jrk()
                    a1 = h1.init_dt(0.0)
                    s1, t1a, _ = h1.run(a1, 1e9, max_steps=k0, trace_cap=k0)
                    s1, t1b, _ = h1.run(s1.atime, 1e9, max_steps=args.integration_steps,
                                        trace_cap=args.integration_steps, itime0=s1.itime)
                    same = bool(np.array_equal(t1a, tn0) and np.array_equal(t1b, tn) and s1.itime == st.itime
                                and s1.atime == st.atime)
                    h1.close()
                same = gather(same)[0]
                check("window_schedule_equals_one_gpu_run", same)

    # ---- the metric's second half: N = 256k integrated to t = 1 N-body time unit ---------------------------------
    nbu = None
    if not args.no_nbu:
        h.upload_particles(r4, v4)
        h.init_acc_jrk()
        at = h.init_dt(0.0)
        e0 = h.get_energy()
        lb = h.launch_count()
        x0 = h.exchange_stats()
        p0 = h.persist_stats()
        barrier()
        st, tnn, _, tun = h.run(at, args.nbu_tend, trace_cap=400000, trace_us=True)
        h.sync()
        barrier()
        wall = max_over_ranks(st.wall_ms) * 1e-3
        e1 = h.get_energy()
        x1 = h.exchange_stats()
        p1 = h.persist_stats()
        small = tnn <= 512
        nbu = {"workload": f"Plummer N={n} equal mass, eps=1e-4, eta=0.01, block-timestep integration t=0..{args.nbu_tend:g} "
                           f"through gdy_run (the reference's while loop, Hermite4CPU_integration.cpp:81-124)",
               "wall_s": wall, "wall_s_per_nbu": wall / st.itime if st.itime > 0 else None, "t_reached": st.itime,
               "block_steps": int(st.iterations), "sum_nact": int(st.nsteps), "pairs": int(st.interactions),
               "gflops": st.interactions / wall * FLOP_PER_PAIR / 1e9,
               "frac_fp64_peak": st.interactions / wall * FLOP_PER_PAIR / 1e12 / (peak_tf * world),
               "e0": e0[0], "e1": e1[0], "cum_rel_energy_error": abs((e1[0] - e0[0]) / e0[0]),
               "block_steps_le_512_active": int(small.sum()), "wall_s_in_block_steps_le_512_active": float(tun[small].sum()) * 1e-6,
               "mean_us_per_block_step_le_512_active": float(tun[small].mean()) if small.any() else None,
               "median_us_per_block_step_nact_le_32": float(np.median(tun[tnn <= 32])) if (tnn <= 32).any() else None,
               "wall_s_by_nact_bin": {f"{lo + 1}..{hi}": {"block_steps": int(((tnn > lo) & (tnn <= hi)).sum()),
                                                          "wall_s": float(tun[(tnn > lo) & (tnn <= hi)].sum()) * 1e-6,
                                                          "pairs": float(tnn[(tnn > lo) & (tnn <= hi)].astype(np.float64).sum()) * n}
                                      for lo, hi in ((0, 16), (16, 512), (512, 4096), (4096, 16384), (16384, 65536), (65536, n))},
               "persistent_launches": p1[0] - p0[0], "persistent_block_steps": p1[1] - p0[1],
               "launches": h.launch_count() - lb, "exchanges_peer_memory": x1[0] - x0[0], "exchanges_nccl": x1[1] - x0[1]}
        if not args.no_checks:
            es = gather((e0[0], e1[0], int(st.iterations), int(st.nsteps)))
            check("nbu_energy_and_schedule_identical_on_all_ranks", len(set(es)) == 1)

    # ---- the north-star target config: full-active sweep at N = 1M ------------------------------------------------
    sweep_1m = None
    if not args.no_sweep_1m:
        from gravidy_b200.plummer import plummer, default_seed
        n1 = 1048576
        m1, r1, v1 = plummer(n1, seed=default_seed(n1))
        h.close()
        h = None
        hm = make_ctx(m1, r1, v1)
        hm.init_acc_jrk()
        hm.sync()
        one_step(hm)
        ms1 = [one_step(hm) for _ in range(3)]
        t1 = max_over_ranks(float(np.mean(ms1))) * 1e-3
        g1 = float(n1) * n1 / t1 * FLOP_PER_PAIR / 1e9
        sweep_1m = {"workload": f"full-active acc+jerk sweep, Plummer N={n1}, eps=1e-4, self pair counted, j-sharded x{world}",
                    "ms_per_sweep": t1 * 1e3, "value": g1, "unit": UNIT, "frac_fp64_peak": g1 / 1e3 / (peak_tf * world),
                    "reps": 3, "l2": "flushed between sweeps"}
        if not args.no_checks and world > 1:
            hs = gather(_sha(hm.download_forces()))
            check("sweep_1m_forces_identical_on_all_ranks", len(set(hs)) == 1, hs[0][:16])
        hm.close()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        crate, info = cpu_force_rate(m, r, v, args.cpu_seconds)
        cpu = {"value": crate * FLOP_PER_PAIR / 1e9, "unit": UNIT, "cores": info["cores"],
               "physical_cores": physical_cores(), "kind": info["kind"],
               "sample": f"update_acc_jrk-style sweep of {info['nact']} of {n} i-particles x {n} j "
                         f"({info['seconds']:.1f} s), scaled to the full N^2"}
        if integ is not None and args.cpu_loop_steps > 0:
            # the same first block steps of the N-body loop on the host cores: the reference's own loop body
            # (find_particles_to_move .. next_integration_time) from the state the GPU run started from
            from oracle import oracle as O
            o = O.best_oracle(m, r, v)
            o.set_threads(info["cores"])
            o.set_state(f6=start["f6"], t=start["t"], dt=start["dt"])
            # bounded: as many of the first block steps as the CPU does in about 2 x --cpu-seconds
            cum = np.cumsum(loop_head["nact"].astype(np.float64) * n)
            k0 = int(min(args.cpu_loop_steps, max(1, np.searchsorted(cum, crate * 2.0 * args.cpu_seconds))))
            loop_head["block_steps"] = k0
            loop_head["pairs"] = int(cum[k0 - 1])
            loop_head["gpu_wall_s"] = float(loop_head["us"][:k0].sum()) * 1e-6
            t0 = time.perf_counter()
            so, tno, _ = o.run(loop_head["atime0"], 1e9, max_steps=k0, trace_cap=k0)
            cpu_wall = time.perf_counter() - t0
            cpu["loop"] = {"workload": f"block-timestep loop, Plummer N={n}, block steps 0..{k0}",
                           "block_steps": int(so.iterations), "pairs": int(so.interactions),
                           "cpu_wall_s": cpu_wall, "gpu_wall_s": loop_head["gpu_wall_s"],
                           "cpu_gflops": so.interactions * FLOP_PER_PAIR / cpu_wall / 1e9,
                           "gpu_gflops": loop_head["pairs"] * FLOP_PER_PAIR / loop_head["gpu_wall_s"] / 1e9,
                           "gpu_over_cpu_wall": cpu_wall / loop_head["gpu_wall_s"],
                           "schedule_identical": bool(np.array_equal(tno, loop_head["nact"][:k0])
                                                      and int(so.interactions) == loop_head["pairs"]),
                           "kind": o.kind, "cores": info["cores"],
                           "note": f"the ratio is against the {info['cores']} host cores of THIS box; a box with twice the cores halves it"}
            if nbu is not None:
                cpu["loop"]["nbu_extrapolated"] = {
                    "cpu_wall_s": nbu["pairs"] / (so.interactions / cpu_wall), "gpu_wall_s": nbu["wall_s"],
                    "gpu_over_cpu_wall": nbu["pairs"] / (so.interactions / cpu_wall) / nbu["wall_s"],
                    "how": "the GPU run's pair count to t = 1 divided by the CPU loop's measured pairs/s"}
            check("cpu_loop_schedule_identical", cpu["loop"]["schedule_identical"])
            o.close()

    if rank == 0:
        line = {
            "metric": METRIC, "value": gflops, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(n), "n": n,
                       "sharding": f"j-sharded x{world}, NCCL allreduce of partial forces" if world > 1 else "1 GPU",
                       "l2": "flushed between timed iterations (256 MiB write)"},
            "pairs_per_s": rate,
            "frac_fp64_peak": gflops / 1e3 / (peak_tf * world),
            "roofline": {"bound": "fp64", "achieved": kern_tf, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": kern_tf / peak_tf,
                         "traffic": NCU_DRAM_BYTES_PER_LAUNCH if (n == 262144 and world == 1) else None,
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full "
                                           "capture of this kernel at N=262144 on 1 GPU "
                                           "(profiles/r02_force_warp_ncu.md): 14.89 MB + 2.65 MB; "
                                           "algorithmic 14.7 MB in + 12.6 MB out",
                         "kernel": "k_force_warp<IPT=2,WI=1,WJ=4,NST=3> (warp_i64, 24 waves of CTAs)",
                         "peak_source": "DFMA micro-benchmark on this GPU (gdy_measure_fp64_peak); "
                                        "MEASURED_PEAKS.json has no FP64 entry",
                         "peak_equiv_sm_mhz_at_64_dfma_per_clk_per_sm": peak_mhz,
                         "nominal_peak_at_max_clock": 2.0 * 64 * 148 * 1965e6 / 1e12,
                         "flop_convention": "60 flop/pair; the kernel issues 31 FP64 instructions/pair"},
            "roofline_hbm": roofline_hbm, "roofline_energy": roofline_energy,
            "e2e": e2e, "gpu_launches": int(launches), "clocks": sampler.summary(),
            "integration": integ, "nbu": nbu, "sweep_1m": sweep_1m,
            "checks": {"ok": not failed, "failed": failed, "items": checks} if not args.no_checks else None,
            "cpu_baseline": cpu, "wall_timed_s": wall_timed,
        }
        print(json.dumps(line), flush=True)
    barrier()
    if h is not None:
        h.close()
    if world > 1:
        dist.destroy_process_group()
    if failed:
        sys.stderr.write("bench.py: CHECKS FAILED: " + ", ".join(failed) + "\n")
        raise SystemExit(3)


def main():
    args = parse_args()
    from gravidy_b200.plummer import plummer, default_seed
    m, r, v = plummer(args.n, seed=default_seed(args.n))
    if args.impl == "reference":
        run_reference(args, m, r, v)
    else:
        run_b200(args, m, r, v)


if __name__ == "__main__":
    main()
