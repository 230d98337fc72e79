#!/bin/bash
# GPU round r1n: fused plane kernel - parity, A/B against the three-kernel path, warp / boxes-per-CTA sweeps, bench
T=r1n; O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_tests.log
for pl in 0 1; do
  echo "### QEB_PLANE=$pl" >> $O/${T}_sweep.log
  QEB_PLANE=$pl python tools/vloc_sweep.py --bands 240 --batches 4,8,15,30 >> $O/${T}_sweep.log 2>&1
done
echo "### plane: warps / boxes per CTA (batch 15)" >> $O/${T}_sweep.log
python tools/vloc_sweep.py --bands 240 --batches 15 \
   --env "QEB_NWP=8;QEB_NWP=10;QEB_NWP=12;QEB_NWP=16;QEB_PLANE_BPC=1;QEB_PLANE_BPC=2;QEB_PLANE_BPC=3;QEB_PLANE_BPC=5" >> $O/${T}_sweep.log 2>&1
python bench.py --steps 5 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench_exit=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${T}_launches_b15.csv \
    python tools/prof_run.py --bands 60 --calls 2 --batch 15 > $O/${T}_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_vloc_plane -s 1 -c 1 -f -o $O/${T}_plane_b15 \
    python tools/prof_run.py --bands 60 --calls 1 --batch 15 >> $O/${T}_ncu.log 2>&1
tail -3 $O/${T}_tests.log; cat $O/${T}_sweep.log; cat $O/${T}_bench.json | cut -c1-400; tail -2 $O/${T}_bench.err
