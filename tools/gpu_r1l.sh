#!/bin/bash
# GPU round r1l: parity tests (incl. s_psi), bench, launch list at the bench's batch size, z/x/y warp + TZ sweeps
T=r1l; O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_tests.log
python bench.py --steps 5 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench_exit=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches_b15.csv \
    python tools/prof_run.py --bands 60 --calls 2 --batch 15 > $O/${T}_ncu.log 2>&1
for tz in 16 32; do
  echo "### TZ=$tz" >> $O/${T}_sweep.log
  QEB_TZ=$tz python tools/vloc_sweep.py --bands 240 --batches 15 \
     --env "QEB_NWZ=4;QEB_NWZ=5;QEB_NWZ=8;QEB_NWZ=10;QEB_NWX=5;QEB_NWX=10;QEB_NWY=4;QEB_NWY=8" >> $O/${T}_sweep.log 2>&1
done
tail -3 $O/${T}_tests.log; cat $O/${T}_bench.json; tail -2 $O/${T}_bench.err; cat $O/${T}_sweep.log
