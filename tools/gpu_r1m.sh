#!/bin/bash
# GPU round r1m: V-prefetch A/B, TY alignment, ncu --set full of the five V_loc kernels at batch 15
T=r1m; O=gpurun_out; mkdir -p $O
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "vloc_psi or h_psi" > $O/${T}_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_tests.log
for lib in "" q-e_b200/libqeb200_novpf.so; do
 for ty in 29 32; do
  echo "### lib=${lib:-default} TY=$ty" >> $O/${T}_sweep.log
  QEB_LIB=$lib QEB_TY=$ty python tools/vloc_sweep.py --bands 240 --batches 4,8,15 >> $O/${T}_sweep.log 2>&1
 done
done
ncu --set full --clock-control none --import-source on -k regex:k_vloc -s 5 -c 5 -f -o $O/${T}_vloc_b15 \
    python tools/prof_run.py --bands 60 --calls 1 --batch 15 > $O/${T}_ncu.log 2>&1
tail -3 $O/${T}_tests.log; cat $O/${T}_sweep.log; tail -2 $O/${T}_ncu.log
