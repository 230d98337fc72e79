#!/bin/bash
# One GPU-box round, parameterised: tools/gpu_round.sh TAG STAGE [STAGE ...]   (outputs under gpurun_out/TAG_*)
#   tests      python -m pytest tests -m gpu            smoke     __graft_entry__.smoke()
#   bench      python bench.py (default line)           ref       python bench.py --impl reference
#   extras     tools/bench_extras.py                    launches  ncu launch list of bench.py (1 step)
#   ncu_vloc   ncu --set full of the V_loc kernels      ncu_gemm  ncu --set full of the projector GEMMs at m=800
#   dist       multi-GPU parity tests (needs gpurun --gpus N)
#   scale      bench.py under torchrun on all GPUs of the box (k-pools + strong-scaling modes)
#   bench:<workload>  bench.py --workload <workload> (si2 al4 h2o64 si512)
#   sweep      tools/vloc_sweep.py with SWEEP_ENV / SWEEP_BATCHES / SWEEP_BANDS
# Every stage runs under its own timeout; a failed stage does not stop the round. ncu stages run only after the same
# command has exited 0 without ncu.
T=${1:-rX}; shift
O=gpurun_out
mkdir -p $O
cd "$(dirname "$0")/.."
NG=$(python -c "import torch;print(torch.cuda.device_count())" 2>/dev/null || echo 1)
for S in "$@"; do
  case $S in
    tests)
      timeout 1500 python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests rc=$?" | tee -a $O/${T}_tests.log
      tail -4 $O/${T}_tests.log;;
    smoke)
      timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${T}_smoke.log;;
    bench)
      timeout 900 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
      cut -c1-600 $O/${T}_bench.json;;
    bench:*)
      W=${S#bench:}
      timeout 900 python bench.py --workload $W --no-cpu-baseline > $O/${T}_bench_$W.json 2> $O/${T}_bench_$W.err; echo "bench $W rc=$?"
      cut -c1-300 $O/${T}_bench_$W.json;;
    ref)
      timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref rc=$?"
      cut -c1-300 $O/${T}_bench_ref.json;;
    extras)
      timeout 900 python tools/bench_extras.py > $O/${T}_extras.jsonl 2> $O/${T}_extras.err; echo "extras rc=$?"; tail -3 $O/${T}_extras.jsonl;;
    launches)
      timeout 300 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-parity --no-library-bar > /dev/null 2>&1 && \
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${T}_launches.csv \
        python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-parity --no-library-bar > $O/${T}_ncu_launches.log 2>&1
      echo "launches rc=$?"; python tools/launch_summary.py $O/${T}_launches.csv $O/${T}_launches_summary.csv;;
    ncu_vloc)
      timeout 300 python tools/prof_run.py --bands 40 --calls 2 > /dev/null 2>&1 && \
      timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_vloc -s 6 -c 6 -f -o $O/${T}_vloc \
        python tools/prof_run.py --bands 40 --calls 2 > $O/${T}_ncu_vloc.log 2>&1
      echo "ncu_vloc rc=$?"; python tools/ncu_summary.py $O/${T}_vloc.ncu-rep $O/${T}_vloc_ncu_full.csv;;
    ncu_gemm)
      timeout 300 python tools/prof_run.py --bands 800 --calls 1 > /dev/null 2>&1 && \
      timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:k_calbec3m|k_addvus3m' -c 2 -f -o $O/${T}_gemm \
        python tools/prof_run.py --bands 800 --calls 1 > $O/${T}_ncu_gemm.log 2>&1
      echo "ncu_gemm rc=$?"; python tools/ncu_summary.py $O/${T}_gemm.ncu-rep $O/${T}_gemm_ncu_full.csv;;
    dist)
      timeout 1500 python -m pytest tests/test_gpu_dist.py -m gpu -x -q > $O/${T}_dist_tests.log 2>&1; echo "dist rc=$?" | tee -a $O/${T}_dist_tests.log
      tail -4 $O/${T}_dist_tests.log;;
    scale)
      for N in $(seq 1 $NG | awk -v n=$NG '$1==1||$1==2||$1==4||$1==8'); do
        if [ $N = 1 ]; then CMD="python bench.py --gpus 1"; else
          CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N"; fi
        timeout 1200 $CMD --no-cpu-baseline > $O/${T}_n${N}.json 2> $O/${T}_n${N}.err; echo "scale N=$N rc=$?"; cut -c1-250 $O/${T}_n${N}.json
      done;;
    sweep)  # V_loc parameter sweep: SWEEP_ENV='A=1;A=2,B=3' SWEEP_BATCHES=20,30 SWEEP_BANDS=400 [SWEEP_WORKLOAD=ausurf112]
      timeout 900 python tools/vloc_sweep.py --workload ${SWEEP_WORKLOAD:-ausurf112} --bands ${SWEEP_BANDS:-400} \
        --batches ${SWEEP_BATCHES:-20} --env "${SWEEP_ENV}" > $O/${T}_sweep.log 2>&1; echo "sweep rc=$?"; cat $O/${T}_sweep.log;;
    *) echo "unknown stage $S";;
  esac
done
