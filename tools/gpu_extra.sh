#!/bin/bash
# extra single-GPU bench lines (other workloads, R&G path on one rank); usage: tools/gpu_extra.sh TAG
T=${1:-rX}; O=gpurun_out; mkdir -p $O
for w in si512 h2o64; do
  python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > $O/${T}_bench_$w.json 2> $O/${T}_bench_$w.err; echo "$w exit=$?"
done
python bench.py --workload si512 --mode rg --steps 3 --warmup 3 --no-cpu-baseline > $O/${T}_bench_si512_rg1.json 2> $O/${T}_bench_si512_rg1.err; echo "rg1 exit=$?"
for w in si2 al4; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/${T}_bench_$w.json 2> $O/${T}_bench_$w.err; echo "$w exit=$?"
done
tail -n 3 $O/${T}_bench_*.err
cat $O/${T}_bench_si512.json $O/${T}_bench_si512_rg1.json $O/${T}_bench_h2o64.json $O/${T}_bench_si2.json $O/${T}_bench_al4.json
