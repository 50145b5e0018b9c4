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
NCU_DRAM_BYTES_PER_LAUNCH = 14995712 + 2934272   # profiles/r01_force_warp_ncu_v12.md


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
    return ap.parse_args()


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
    rates, info = [], None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        rate, info = cpu_force_rate(m, r, v, per_step)
        rates.append(rate)
    wall = time.perf_counter() - t0
    rate = float(np.mean(rates))
    gflops = rate * FLOP_PER_PAIR / 1e9
    sample = (f"update_acc_jrk-style sweep of {info['nact']} of {n} i-particles x {n} j per step, "
              f"scaled to the full N^2")
    line = {
        "impl": "reference", "metric": METRIC, "value": gflops, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": n * float(n) / rate * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"full-active acc+jerk sweep, Plummer N={n}, eps=1e-4", "n": n},
        "pairs_per_s": rate,
        "cpu_baseline": {"value": gflops, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                         "sample": sample},
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
    uid = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        t = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(Hermite4B200.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        uid = bytes(t.cpu().numpy().tobytes())

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

    n = len(m)
    h = Hermite4B200(m, r, v, softening=1e-4, eta=0.01, device=local, rank=rank, world=world, nccl_uid=uid)
    if world > 1:
        # peer memory for the fused NVLink exchange of the persistent small-block kernel
        hs = [None] * world
        dist.all_gather_object(hs, h.peer_export())
        h.peer_attach(hs)
        dist.barrier()
    h.init_acc_jrk()
    h.sync()
    peak_tf, peak_mhz = h.measure_fp64_peak()
    pairs = float(n) * float(n)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def one_step():
        flush.zero_()
        barrier()
        return h.force_sweep_full(1)

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

    # ---- the memory-bound kernel of the step: the predictor (120 B read + 48 B written per particle) ----
    pred_ms = []
    for _ in range(3 + 5):
        flush.zero_()
        torch.cuda.synchronize()
        pred_ms.append(h.predict_timed(0.0))
    pred_ms = float(np.mean(pred_ms[3:]))
    hbm_peak, hbm_src = 6551.4, "fallback: the B200 copy bandwidth of MEASURED_PEAKS.json on this pool (file absent)"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            hbm_peak, hbm_src = float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (copy, read+write bytes)"
    except Exception:
        pass
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

    # ---- e2e: host buffers through the C ABI, copies inside the timed region --------------------------
    r4 = torch.empty((n, 4), dtype=torch.float64).pin_memory().numpy()
    v4 = torch.empty((n, 4), dtype=torch.float64).pin_memory().numpy()
    f6 = torch.empty((n, 6), dtype=torch.float64).pin_memory().numpy()
    r4[:, :3] = r; r4[:, 3] = m; v4[:, :3] = v; v4[:, 3] = 0.0

    def e2e_step():
        h.upload_particles(r4, v4)
        h.init_acc_jrk()
        h.download_forces(out=f6)

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
    e2e = {"value": pairs / e2e_s * FLOP_PER_PAIR / 1e9, "unit": UNIT,
           "h2d_bytes_per_step": int(r4.nbytes + v4.nbytes), "d2h_bytes_per_step": int(f6.nbytes),
           "ms_per_step": e2e_s * 1e3,
           "call": "gdy_upload_particles + gdy_init_acc_jrk + gdy_download_forces (pinned host buffers)"}

    # ---- a window of the block-timestep integration (small active blocks) -----------------------------
    integ = None
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
        barrier()
        st, tn, _, tu = h.run(st.atime, 1e9, max_steps=args.integration_steps,
                              trace_cap=args.integration_steps, itime0=st.itime, trace_us=True)
        barrier()
        wall = max_over_ranks(st.wall_ms) * 1e-3
        integ = {"workload": f"block-timestep integration, Plummer N={n}, block steps {k0}..{k0 + st.iterations}",
                 "block_steps": int(st.iterations), "sum_nact": int(st.nsteps),
                 "mean_nact": float(st.nsteps) / max(int(st.iterations), 1),
                 "median_nact": float(np.median(tn)) if len(tn) else None,
                 "pairs": int(st.interactions), "wall_s": wall,
                 "us_per_block_step": wall / max(int(st.iterations), 1) * 1e6,
                 "gflops": st.interactions / wall * FLOP_PER_PAIR / 1e9,
                 "launches": h.launch_count() - l1, "t_reached": st.itime,
                 "us_per_block_step_nact_le_32": float(np.median(tu[tn <= 32])) if (tn <= 32).any() else None}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        crate, info = cpu_force_rate(m, r, v, args.cpu_seconds)
        cpu = {"value": crate * FLOP_PER_PAIR / 1e9, "unit": UNIT, "cores": info["cores"],
               "kind": info["kind"],
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
                           "kind": o.kind, "cores": info["cores"]}
            o.close()

    if rank == 0:
        line = {
            "metric": METRIC, "value": gflops, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"full-active acc+jerk sweep (init_acc_jrk force evaluation), Plummer "
                                   f"N={n}, eps=1e-4, self pair counted", "n": n,
                       "sharding": f"j-sharded x{world}, NCCL allreduce of partial forces" if world > 1 else "1 GPU",
                       "l2": "flushed between timed iterations (256 MiB write)"},
            "pairs_per_s": rate,
            "frac_fp64_peak": gflops / 1e3 / (peak_tf * world),
            "roofline": {"bound": "fp64", "achieved": kern_tf, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": kern_tf / peak_tf,
                         "traffic": NCU_DRAM_BYTES_PER_LAUNCH if (n == 262144 and world == 1) else None,
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full "
                                           "capture of this kernel at N=262144 on 1 GPU "
                                           "(profiles/r01_force_warp_ncu_v12.md): 15.00 MB + 2.93 MB; "
                                           "algorithmic 14.7 MB in + 12.6 MB out",
                         "kernel": "k_force_warp<IPT=2,WI=1,WJ=4,NST=3> (warp_i64, 24 waves of CTAs)",
                         "peak_source": "DFMA micro-benchmark on this GPU (gdy_measure_fp64_peak); "
                                        "MEASURED_PEAKS.json has no FP64 entry",
                         "peak_equiv_sm_mhz_at_64_dfma_per_clk_per_sm": peak_mhz,
                         "nominal_peak_at_max_clock": 2.0 * 64 * 148 * 1965e6 / 1e12,
                         "flop_convention": "60 flop/pair; the kernel issues 31 FP64 instructions/pair"},
            "roofline_hbm": roofline_hbm,
            "e2e": e2e, "gpu_launches": int(launches), "clocks": sampler.summary(),
            "integration": integ, "cpu_baseline": cpu, "wall_timed_s": wall_timed,
        }
        print(json.dumps(line), flush=True)
    barrier()
    h.close()
    if world > 1:
        dist.destroy_process_group()


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
