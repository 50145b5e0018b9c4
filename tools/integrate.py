#!/usr/bin/env python
"""Block-timestep integration of a seeded Plummer model through the fused driver (gdy_run): the
BASELINE.json configs 2-4 (N=16k equal mass, N=64k Salpeter, N=256k j-sharded over 1/2/4/8 GPUs).

Prints ONE JSON line: block steps, sum of nact, pairs, wall seconds of the `while` loop (SURVEY.md 8d:
init_acc_jrk, the energy diagnostic and nbody_attributes are outside the window on both sides),
GFLOP/s at 60 flop/pair, the energy error at the end, and the per-nact-bin step times.

    python tools/integrate.py --particles 16384 --tend 1
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/integrate.py --particles 262144 --tend 1
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gravidy_b200.hermite4 import Hermite4B200  # noqa: E402
from gravidy_b200.plummer import plummer, default_seed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", "-N", dest="n", type=int, default=16384)
    ap.add_argument("--tend", "-T", dest="t", type=float, default=1.0)
    ap.add_argument("--spectrum", default="equal")
    ap.add_argument("--softening", type=float, default=1e-4)
    ap.add_argument("--eta", type=float, default=0.01)
    ap.add_argument("--max-steps", type=int, default=0)
    ap.add_argument("--trace", type=int, default=400000)
    ap.add_argument("--out", default="")
    ap.add_argument("--dump-trace", default="", help="npz of the per-step trace (nact, ITIME, us) of rank 0")
    ap.add_argument("--no-peers", action="store_true", help="sharded run without the fused NVLink exchange")
    ap.add_argument("--rank-prof", default="", help="every rank writes its per-nact-bin step times and the persistent "
                                                   "kernel's phase profile (GDY_PERSIST_PROF=1) to <prefix>.rank<r>.txt")
    a = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    uid = None
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
        dist.init_process_group("nccl", device_id=dev)
        t = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(Hermite4B200.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        uid = bytes(t.cpu().numpy().tobytes())

    m, r, v = plummer(a.n, seed=default_seed(a.n), mass_spectrum=a.spectrum)
    h = Hermite4B200(m, r, v, softening=a.softening, eta=a.eta, device=local, rank=rank, world=world,
                     nccl_uid=uid)
    if world > 1 and not a.no_peers:
        hs = [None] * world
        dist.all_gather_object(hs, h.peer_export())
        h.peer_attach(hs)
        dist.barrier()
    t0 = time.perf_counter()
    h.init_acc_jrk()
    h.sync()
    t_init = time.perf_counter() - t0
    at = h.init_dt(0.0)
    t0 = time.perf_counter()
    e0 = h.get_energy()
    t_energy = time.perf_counter() - t0
    if dist is not None:
        dist.barrier()
    st, tn, ti, tu = h.run(at, a.t, max_steps=a.max_steps, trace_cap=a.trace, trace_us=True)
    h.sync()
    wall = st.wall_ms * 1e-3
    if dist is not None:
        import torch
        tt = torch.tensor([wall], dtype=torch.float64, device=torch.device("cuda", local))
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        wall = float(tt.item())
    e1 = h.get_energy()
    bins = []
    edges = [0, 1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 4096, 16384, 65536, 1 << 30]
    for lo, hi in zip(edges[:-1], edges[1:]):
        sel = (tn > lo) & (tn <= hi)
        if sel.any():
            bins.append({"nact": [lo, hi], "steps": int(sel.sum()), "median_us": float(np.median(tu[sel])),
                         "share_of_wall": float(tu[sel].sum() / tu.sum())})
    line = {
        "tool": "integrate", "n": a.n, "n_gpus": world, "spectrum": a.spectrum, "softening": a.softening,
        "eta": a.eta, "t_end": a.t, "t_reached": st.itime, "block_steps": int(st.iterations),
        "sum_nact": int(st.nsteps), "pairs": int(st.interactions), "wall_s": wall,
        "wall_s_per_nbu": wall / st.itime if st.itime > 0 else None,
        "us_per_block_step": wall / max(int(st.iterations), 1) * 1e6,
        "pairs_per_s": st.interactions / wall, "gflops_60": st.interactions / wall * 60 / 1e9,
        "force_kernel_s": st.force_ms * 1e-3,
        "e0": e0[0], "e1": e1[0], "cum_rel_energy_error": abs((e1[0] - e0[0]) / e0[0]),
        "init_acc_jrk_s": t_init, "energy_s": t_energy, "launches": h.launch_count(), "bins": bins,
    }
    line["persist_launches"], line["persist_steps"] = h.persist_stats()
    if a.rank_prof:
        prof = h.persist_profile()
        ps = max(line["persist_steps"], 1)
        names = ["P1a", "P1b", "bar1", "P2", "bar2", "P3", "exit", "-", "b.stage", "b.sync", "b.pairs", "b.foldsync"]
        if os.environ.get("GDY_PERSIST_V", "2") != "1":
            names = ["load", "A.publish", "B.predict+scan", "P1.wait_i", "P1.pairs+publish", "P2", "P3.wait", "P3.lists",
                     "exit", "-", "-", "-", "xchg.wait_total", "xchg.count"]
        with open(f"{a.rank_prof}.rank{rank}.txt", "w") as f:
            f.write(f"rank {rank}/{world} wall {st.wall_ms:.1f} ms steps {st.iterations} persist_steps {ps}\n")
            if len(prof) > 14 and prof[14]:
                ps = prof[14]   # the timed steps (GDY_PERSIST_PROF_NACT: only steps with at most that many active particles)
                f.write(f"timed steps {ps} (nact <= {os.environ.get('GDY_PERSIST_PROF_NACT', 'any')})\n")
            f.write("ns/step by phase (CTA 0): " + "  ".join(f"{nm} {v / ps:.0f}" for nm, v in zip(names, prof) if v) + "\n")
            if len(prof) > 13 and prof[13]:
                f.write(f"exchange wait per owned particle: {prof[12] / prof[13]:.0f} ns over {prof[13]} waits\n")
            for b in bins:
                f.write(f"nact ({b['nact'][0]},{b['nact'][1]}] steps {b['steps']} median_us {b['median_us']:.1f} "
                        f"total_ms {b['share_of_wall'] * tu.sum() / 1e3:.2f}\n")
    if rank == 0 and a.dump_trace:
        np.savez_compressed(a.dump_trace, nact=tn, itime=ti, us=tu)
    if rank == 0:
        s = json.dumps(line)
        print(s, flush=True)
        if a.out:
            with open(a.out, "w") as f:
                f.write(s + "\n")
    if dist is not None:
        dist.barrier()
    h.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
