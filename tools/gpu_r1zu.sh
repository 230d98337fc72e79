#!/bin/bash
# ncu --set full of the V_loc kernels at the final batch (20 boxes) and of the spinor x pass
T=r1zu; O=gpurun_out; mkdir -p $O
cd /root/repo
python tools/prof_run.py --bands 80 --calls 1 > $O/${T}_run.log 2>&1 || exit 1
python tools/prof_run.py --bands 80 --calls 1 --mode nc >> $O/${T}_run.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_vloc -s 3 -c 3 -f -o $O/${T}_vloc_b20 \
    python tools/prof_run.py --bands 80 --calls 1 > $O/${T}_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_vloc_xspin -s 1 -c 1 -f -o $O/${T}_xspin \
    python tools/prof_run.py --bands 80 --calls 1 --mode nc >> $O/${T}_ncu.log 2>&1
tail -3 $O/${T}_ncu.log; ls -la $O/${T}_*
