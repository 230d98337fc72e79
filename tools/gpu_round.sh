#!/bin/bash
# One GPU-box round: parity tests, bench (ours + reference arm), ncu launch list and --set full captures.
# usage: tools/gpu_round.sh TAG [full]   (outputs under gpurun_out/TAG_*)
T=${1:-rX}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_tests.log
python bench.py --steps 5 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench_exit=$?"
if [ "$2" = "full" ]; then
  python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${T}_launches.csv \
      python tools/prof_run.py --bands 64 --calls 2 > $O/${T}_ncu.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:k_vloc -s 80 -c 5 -f -o $O/${T}_vloc \
      python tools/prof_run.py --bands 64 --calls 2 >> $O/${T}_ncu.log 2>&1
  ncu --set full --clock-control none --import-source on -k 'regex:^k_calbec$|^k_addvus$' -s 2 -c 2 -f -o $O/${T}_gemm \
      python tools/prof_run.py --bands 64 --calls 2 >> $O/${T}_ncu.log 2>&1
fi
tail -3 $O/${T}_tests.log; cat $O/${T}_bench.json; tail -2 $O/${T}_bench.err
