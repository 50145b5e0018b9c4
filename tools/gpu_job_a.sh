#!/bin/bash
# One gpurun call: GPU tests, the bench line (both arms), ncu launch list + full captures, the N sweep
# and the BASELINE configs 2-4 on one GPU.  Everything lands in gpurun_out/.
#   gpurun --timeout 2400 -- bash tools/gpu_job_a.sh
set -u
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi.txt 2>&1

echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 $O/pytest_gpu.log

echo "== bench (default)"; timeout 600 python bench.py > $O/bench_1gpu.json 2> $O/bench_1gpu.err; echo "rc=$?"; cat $O/bench_1gpu.json
echo "== bench --impl reference"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "rc=$?"; cat $O/bench_ref.json

echo "== N sweep (full-active)"
for n in 65536 131072 524288 1048576; do
  timeout 600 python bench.py --n $n --steps 3 --warmup 3 --no-cpu-baseline --no-integration > $O/bench_n$n.json 2>> $O/bench_sweep.err
  python - <<EOF
import json
d = json.load(open("$O/bench_n$n.json"))
print($n, "GFLOP/s", round(d["value"]), "frac", round(d["roofline"]["frac"], 4), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]))
EOF
done

echo "== integrations"
timeout 300 python tools/integrate.py --particles 16384 --tend 1 --out $O/integrate_n16k.json > /dev/null 2> $O/integrate.err; echo "rc=$?"
timeout 600 python tools/integrate.py --particles 65536 --tend 0.25 --spectrum salpeter --out $O/integrate_n64k_salpeter.json > /dev/null 2>> $O/integrate.err; echo "rc=$?"
python tools/step_latency.py --n 262144 --steps 1500 > $O/step_latency_n256k.txt 2>&1
python tools/step_latency.py --n 16384 --steps 3000 > $O/step_latency_n16k.txt 2>&1

echo "== ncu launch list"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --integration-steps 60 > $O/ncu_launches.log 2>&1; echo "rc=$?"

echo "== ncu full: force kernel (full-active)"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_force_warp -s 4 -c 1 -f -o $O/prof_force_warp \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-integration > $O/ncu_force.log 2>&1; echo "rc=$?"
echo "== ncu full: step kernels"
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_predict|k_correct|k_min_time|k_sched|k_force_lane|k_reduce' -s 120 -c 14 -f -o $O/prof_step \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --integration-steps 60 > $O/ncu_step.log 2>&1; echo "rc=$?"

echo "== full N=256k to t=1 (1 GPU)"
timeout 900 python tools/integrate.py --particles 262144 --tend 1 --out $O/integrate_n256k_1gpu.json > /dev/null 2>> $O/integrate.err; echo "rc=$?"
python - <<EOF
import json
for f in ("integrate_n16k", "integrate_n64k_salpeter", "integrate_n256k_1gpu"):
    try:
        d = json.load(open("$O/%s.json" % f))
        print(f, "steps", d["block_steps"], "wall", round(d["wall_s"], 2), "GF", round(d["gflops_60"]), "dE", d["cum_rel_energy_error"])
    except Exception as e:
        print(f, "missing", e)
EOF
