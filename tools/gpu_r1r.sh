#!/bin/bash
# 8-GPU box: distributed parity tests, weak scaling over k-pools at N=2,4,8, band groups and R&G slab FFT at N=8
T=r1r; O=gpurun_out; mkdir -p $O
python -m pytest tests/test_gpu_dist.py -m gpu -x -q > $O/${T}_dist_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_dist_tests.log
run() { # N mode workload tag
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $1 --steps 3 --warmup 3 --mode $2 --workload $3 > $O/${T}_n$1_$4.json 2> $O/${T}_n$1_$4.err
  echo "n$1 $4 exit=$?"
}
python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > $O/${T}_n1_pools.json 2> $O/${T}_n1_pools.err
run 2 pools ausurf112 pools
run 4 pools ausurf112 pools
run 8 pools ausurf112 pools
run 8 bgrp ausurf112 bgrp
run 8 rg si512 rg_si512
run 2 rg si512 rg_si512
tail -3 $O/${T}_dist_tests.log
for f in $O/${T}_n*.json; do echo $f; python - $f <<'P'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(j["n_gpus"], j["config"]["parallelism"], round(j["value"],1), round(j["ms_per_step"],2), j.get("clocks"), j.get("sections_ms"))
except Exception as e: print("bad", e)
P
done
tail -n 3 $O/${T}_n8_*.err
