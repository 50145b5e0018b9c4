#!/bin/bash
# One parameterised GPU job: `gpurun [--gpus N] -- 'bash tools/gpu_job.sh TAG stage [stage ...]'`.
# Every stage writes gpurun_out/<TAG>_<stage>.{log,json,...}; a stage is `name` or `name:arg1:arg2`.
#
#   tests            pytest -m gpu (whole suite)                 tests:<file-or-k-expr>  a subset
#   hooks[:n]        tools/hook_latency.py: the hook loop beside gdy_run, production and lab builds
#   configs          BASELINE configs 2 and 3 to t = 1 on one GPU + the sweep at N = 64k, 128k, 512k
#   p2prof[:n:steps] phase profile of the persistent kernel by active-set size on a small model (cap 2048)
#   smoke            __graft_entry__.build() + smoke()
#   bench[:N]        bench.py on N GPUs (default 1)               benchref  bench.py --impl reference
#   lat:<n>[:spec]   tools/step_latency.py at n particles         integ:<n>:<tend>[:gpus[:extra env]]  tools/integrate.py
#   ncu_force        ncu --set full of the force kernel in bench.py (one capture) + launch list
#   ncu_energy       ncu --set full of k_energy and k_knn_distance at N = 256k
#   lab[:nacts]      every build under gravidy_b200/lab/*.so against the production library: bits and big-block rates
#   sanit            compute-sanitizer memcheck + racecheck of a short persistent-kernel run
#   diag4            the 4-GPU small-step diagnostic of round 2 (phase profile of every rank)
set -u
TAG=$1; shift
O=gpurun_out
mkdir -p $O
export PYTHONUNBUFFERED=1
TR() { np=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + np)) "$@"; }
for st in "$@"; do
  IFS=: read -r name a1 a2 a3 a4 <<< "$st"
  t0=$(date +%s)
  case $name in
    tests)
      if [ -z "${a1:-}" ]; then timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_tests.log 2>&1
      elif [ -f "tests/$a1" ]; then timeout 1500 python -m pytest tests/$a1 -m gpu -x -q > $O/${TAG}_tests.log 2>&1
      else timeout 1500 python -m pytest tests -m gpu -x -q -k "$a1" > $O/${TAG}_tests.log 2>&1; fi
      echo "[$st] rc=$? $(tail -1 $O/${TAG}_tests.log)";;
    bench)
      g=${a1:-1}
      if [ "$g" = 1 ]; then timeout 1500 python bench.py ${a2:-} > $O/${TAG}_bench_1gpu.json 2> $O/${TAG}_bench_1gpu.err
      else timeout 1500 bash -c "$(declare -f TR); TR $g bench.py --gpus $g ${a2:-}" > $O/${TAG}_bench_${g}gpu.json 2> $O/${TAG}_bench_${g}gpu.err; fi
      echo "[$st] rc=$? $(tail -c 600 $O/${TAG}_bench_${g}gpu.json)";;
    smoke)
      timeout 600 python __graft_entry__.py --smoke > $O/${TAG}_smoke.log 2>&1
      echo "[$st] rc=$? $(tail -1 $O/${TAG}_smoke.log)";;
    hooks)
      # hooks[:n]: the hook loop's per-step time, production library and every build under gravidy_b200/lab/
      timeout 300 python tools/hook_latency.py --n ${a1:-262144} > $O/${TAG}_hooks.txt 2>&1
      for L in gravidy_b200/lab/*.so; do [ -f "$L" ] && GDY_LIB=$PWD/$L timeout 300 python tools/hook_latency.py --n ${a1:-262144} >> $O/${TAG}_hooks.txt 2>&1; done
      cat $O/${TAG}_hooks.txt;;
    configs)
      # BASELINE.json configs 2, 3 (N = 16k equal mass, N = 64k Salpeter, to t = 1) and the other sizes of config 5's sweep
      timeout 300 python tools/integrate.py --particles 16384 --tend 1 > $O/${TAG}_integ_n16k.json 2> $O/${TAG}_integ_n16k.err
      timeout 300 python tools/integrate.py --particles 65536 --tend 1 --spectrum salpeter > $O/${TAG}_integ_n64k_salpeter.json 2> $O/${TAG}_integ_n64k_salpeter.err
      for n in 65536 131072 524288; do timeout 300 python tools/sweep_big.py --n $n --nacts $n --variants 2 --reps 3; done > $O/${TAG}_sweep_sizes.txt 2>&1
      python - <<PY
import json
for f in ("$O/${TAG}_integ_n16k.json", "$O/${TAG}_integ_n64k_salpeter.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "wall", round(d["wall_s"], 3), "steps", d["block_steps"], "E", d["cum_rel_energy_error"], "persist", d["persist_steps"], "gflops", round(d.get("gflops", 0)))
    except Exception as e:
        print(f, "no result:", e)
PY
      grep -E "^[0-9]+,[0-9]+,-1" $O/${TAG}_sweep_sizes.txt;;
    p2prof)
      # p2prof:<n>:<steps>: the persistent kernel with the cap at 2048 on a small model (N = 32768 on one GPU gives every CTA the
      # 222-particle slice of an N = 256k run on eight): per-step medians by nact bin, then the phase profile of CTA 0
      # restricted to block steps of at most 64 ... 2048 active particles (differences = cost of the extra i-chunks)
      n=${a1:-32768}; k=${a2:-6000}
      GDY_PERSIST_NACT=2048 timeout 300 python tools/step_latency.py --n $n --steps $k > $O/${TAG}_p2prof.txt 2>&1
      for f in 16 64 128 256 512 1024 2048; do
        echo "== phases of the block steps with nact <= $f" >> $O/${TAG}_p2prof.txt
        GDY_PERSIST_NACT=2048 GDY_PERSIST_PROF=1 GDY_PERSIST_PROF_NACT=$f timeout 300 python tools/step_latency.py --n $n --steps $k 2>&1 | grep -E "ns/step|persistent kernel" >> $O/${TAG}_p2prof.txt
      done
      cat $O/${TAG}_p2prof.txt;;
    benchref)
      timeout 900 python bench.py --impl reference > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
      echo "[$st] rc=$? $(tail -c 300 $O/${TAG}_bench_reference.json)";;
    lat)
      timeout 600 python tools/step_latency.py --n $a1 --spectrum ${a2:-equal} ${a3:+--steps $a3} > $O/${TAG}_lat_n$a1.txt 2>&1
      echo "[$st] rc=$?"; head -30 $O/${TAG}_lat_n$a1.txt
      GDY_PERSIST_PROF=1 timeout 600 python tools/step_latency.py --n $a1 --spectrum ${a2:-equal} ${a3:+--steps $a3} 2>&1 | grep -E "ns/step|median" | head -4;;
    integ)
      g=${a3:-1}
      if [ "$g" = 1 ]; then env ${a4:-X=1} timeout 900 python tools/integrate.py --particles $a1 --tend $a2 --rank-prof $O/${TAG}_integ_n${a1}_${g}gpu > $O/${TAG}_integ_n${a1}_${g}gpu.json 2> $O/${TAG}_integ_n${a1}_${g}gpu.err
      else env ${a4:-X=1} timeout 900 bash -c "$(declare -f TR); TR $g tools/integrate.py --particles $a1 --tend $a2 --rank-prof $O/${TAG}_integ_n${a1}_${g}gpu" > $O/${TAG}_integ_n${a1}_${g}gpu.json 2> $O/${TAG}_integ_n${a1}_${g}gpu.err; fi
      echo "[$st] rc=$? $(python - <<PY
import json
try:
    d = json.loads(open("$O/${TAG}_integ_n${a1}_${g}gpu.json").read().strip().splitlines()[-1])
    print("wall", round(d["wall_s"], 4), "steps", d["block_steps"], "E", d["cum_rel_energy_error"], "persist", d["persist_steps"],
          " ".join("%s:%.1f" % (b["nact"][1], b["median_us"]) for b in d["bins"]))
except Exception as e:
    print("no result:", e)
PY
)";;
    ncu_force)
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_force_warp -c 1 -o $O/${TAG}_force_warp \
        python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-integration --no-extras > $O/${TAG}_ncu_force.log 2>&1
      echo "[$st] rc=$?"
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
        python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > $O/${TAG}_ncu_launches.log 2>&1
      echo "[$st] launches rc=$?";;
    ncu_energy)
      # one ncu --set full capture each of the energy and the k-th-neighbour kernels at N = 256k (FP64-bound)
      cat > /tmp/ncu_energy.py <<'PY'
import sys; sys.path.insert(0, ".")
from gravidy_b200.hermite4 import Hermite4B200
from gravidy_b200.plummer import plummer, default_seed
n = 262144
m, r, v = plummer(n, seed=default_seed(n))
h = Hermite4B200(m, r, v, softening=1e-4, eta=0.01, device=0)
print("energy ms", h.energy_timed()); print("knn ms", h.neighbour_timed(6))
PY
      for K in k_energy k_knn_distance; do
        timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${K}\$" -c 1 -f -o $O/${TAG}_$K \
          python /tmp/ncu_energy.py > $O/${TAG}_ncu_$K.log 2>&1
        echo "[$st] $K rc=$? $(tail -2 $O/${TAG}_ncu_$K.log | tr '\n' ' ')"
      done;;
    ncu_step)
      # ncu --set full of the memory-bound step kernels at N = 256k (predictor, corrector, scheduler passes): the hook
      # sequence of a few block steps with the persistent kernel off; the program runs once WITHOUT ncu first
      cat > /tmp/ncu_step.py <<'PY'
import sys; sys.path.insert(0, ".")
from gravidy_b200.hermite4 import Hermite4B200
from gravidy_b200.plummer import plummer, default_seed
n = 262144
m, r, v = plummer(n, seed=default_seed(n))
h = Hermite4B200(m, r, v, softening=1e-4, eta=0.01, device=0)
h.set_persist(0)
h.init_acc_jrk(); at = h.init_dt(0.0)
for s in range(12):
    nact = h.find_particles_to_move(at)
    h.save_old_acc_jrk(nact); h.predicted_pos_vel(at); h.update_acc_jrk(nact); h.correction_pos_vel(at, nact)
    at = h.next_integration_time()
print("steps ok, last nact", nact, "predict ms", h.predict_timed(at))
PY
      timeout 300 python /tmp/ncu_step.py > $O/${TAG}_ncu_step_plain.log 2>&1 && \
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_predict|k_correct|k_min_time|k_sched_count|k_sched_write|k_save_old" \
        --launch-skip 6 -c 12 -f -o $O/${TAG}_step_kernels python /tmp/ncu_step.py > $O/${TAG}_ncu_step.log 2>&1
      echo "[$st] rc=$? $(tail -2 $O/${TAG}_ncu_step.log | tr '\n' ' ')";;
    hostprog)
      # hostprog:<n>:<tend>:<gpus>[:extra args]  the product binary timed from the outside (tools/host_program_run.py)
      timeout 1500 python tools/host_program_run.py --particles ${a1:-16384} --tend ${a2:-1} --gpus ${a3:-1} ${a4:-} \
        >> $O/${TAG}_hostprog.jsonl 2> $O/${TAG}_hostprog.err
      echo "[$st] rc=$? $(tail -1 $O/${TAG}_hostprog.jsonl | cut -c1-700)";;
    lab)
      # lab[:nacts]: every library build under gravidy_b200/lab/ against the production library: forces bit for bit
      # (tools/lab/force_cmp.sh) and the big-block rates (tools/sweep_big.py)
      P=gravidy_b200/libgravidy_b200.so
      echo "== production" > $O/${TAG}_lab.txt
      timeout 300 python tools/sweep_big.py --n 262144 --nacts ${a1:-1172,8192,262144} --variants 2 --reps 4 >> $O/${TAG}_lab.txt 2>&1
      for L in gravidy_b200/lab/*.so; do
        echo "== $L" >> $O/${TAG}_lab.txt
        timeout 300 bash tools/lab/force_cmp.sh $P $L 4099 65536 >> $O/${TAG}_lab.txt 2>&1
        GDY_LIB=$PWD/$L timeout 300 python tools/sweep_big.py --n 262144 --nacts ${a1:-1172,8192,262144} --variants 2 --reps 4 >> $O/${TAG}_lab.txt 2>&1
      done
      echo "== production again" >> $O/${TAG}_lab.txt
      timeout 300 python tools/sweep_big.py --n 262144 --nacts ${a1:-1172,8192,262144} --variants 2 --reps 4 >> $O/${TAG}_lab.txt 2>&1
      grep -E "^==|differing|^262144" $O/${TAG}_lab.txt;;
    sanit)
      timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_smoke.py > $O/${TAG}_sanit_memcheck.log 2>&1
      echo "[$st] memcheck rc=$? $(tail -2 $O/${TAG}_sanit_memcheck.log | tr '\n' ' ')"
      GDY_SANIT_SHORT=1 timeout 1200 compute-sanitizer --tool racecheck --error-exitcode 9 python tools/sanitize_smoke.py > $O/${TAG}_sanit_racecheck.log 2>&1
      echo "[$st] racecheck rc=$? $(tail -2 $O/${TAG}_sanit_racecheck.log | tr '\n' ' ')";;
    diag4)
      export GDY_PERSIST_PROF=1
      T=0.0078125
      timeout 300 python tools/integrate.py --particles 262144 --tend $T --rank-prof $O/${TAG}_diag_1gpu > $O/${TAG}_diag_1gpu.json 2> $O/${TAG}_diag_1gpu.err
      timeout 300 bash -c "$(declare -f TR); TR 4 tools/integrate.py --particles 262144 --tend $T --rank-prof $O/${TAG}_diag_4gpu_default" > $O/${TAG}_diag_4gpu_default.json 2> $O/${TAG}_diag_4gpu_default.err
      GDY_REPL_NACT=2048 timeout 300 bash -c "$(declare -f TR); TR 4 tools/integrate.py --particles 262144 --tend $T --rank-prof $O/${TAG}_diag_4gpu_repl" > $O/${TAG}_diag_4gpu_repl.json 2> $O/${TAG}_diag_4gpu_repl.err
      for g in 0 1 2 3; do
        CUDA_VISIBLE_DEVICES=$g timeout 300 python tools/integrate.py --particles 262144 --tend $T --rank-prof $O/${TAG}_diag_indep$g > $O/${TAG}_diag_indep$g.json 2> $O/${TAG}_diag_indep$g.err &
      done
      wait; unset GDY_PERSIST_PROF
      tail -n 30 $O/${TAG}_diag_*.txt;;
    exp)
      # exp:<gpus>:<tend>: the persistent kernel's phase profile for tiny steps on every rank, then the persistent cap A/B
      g=${a1:-8}; T=${a2:-0.125}
      GDY_PERSIST_PROF=1 GDY_PERSIST_PROF_NACT=16 timeout 300 bash -c "$(declare -f TR); TR $g tools/integrate.py --particles 262144 --tend $T --rank-prof $O/${TAG}_exp_prof16_${g}gpu" > $O/${TAG}_exp_prof16_${g}gpu.json 2> $O/${TAG}_exp_prof16_${g}gpu.err
      for cap in 512 1024 2048; do
        GDY_PERSIST_NACT=$cap timeout 300 bash -c "$(declare -f TR); TR $g tools/integrate.py --particles 262144 --tend $T --rank-prof $O/${TAG}_exp_cap${cap}_${g}gpu" > $O/${TAG}_exp_cap${cap}_${g}gpu.json 2> $O/${TAG}_exp_cap${cap}_${g}gpu.err
        echo "cap $cap: $(python -c "import json;d=json.loads(open('$O/${TAG}_exp_cap${cap}_${g}gpu.json').read().strip().splitlines()[-1]);print('wall',round(d['wall_s'],4),' '.join('%s:%.0f'%(b['nact'][1],b['median_us']) for b in d['bins']))" 2>&1)"
      done
      cat $O/${TAG}_exp_prof16_${g}gpu.rank0.txt | head -4;;
    *) echo "unknown stage $st";;
  esac
  echo "[$st] $(( $(date +%s) - t0 )) s"
done
