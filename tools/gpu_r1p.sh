#!/bin/bash
# GPU round r1p: low-register radix plans (twice the resident warps) for the z passes and the plane kernel
T=r1p; O=gpurun_out; mkdir -p $O
S="python tools/vloc_sweep.py --bands 240 --batches 15"
echo "### default plans" >> $O/${T}_sweep.log;            $S >> $O/${T}_sweep.log 2>&1
echo "### ZPLAN=8,5,5" >> $O/${T}_sweep.log;              QEB_ZPLAN=8,5,5 $S --env "QEB_NWZ=5;QEB_NWZ=8;QEB_NWZ=10" >> $O/${T}_sweep.log 2>&1
echo "### ZPLAN=8,5,5 TZ=32" >> $O/${T}_sweep.log;        QEB_TZ=32 QEB_ZPLAN=8,5,5 $S --env "QEB_NWZ=5;QEB_NWZ=8;QEB_NWZ=10" >> $O/${T}_sweep.log 2>&1
echo "### XPLAN=5,5,5 (plane 1024 thr)" >> $O/${T}_sweep.log; QEB_XPLAN=5,5,5 $S --env "QEB_NWP=16;QEB_NWP=24;QEB_NWP=32" >> $O/${T}_sweep.log 2>&1
echo "### XPLAN=5,5,5 YPLAN=4,4,4" >> $O/${T}_sweep.log;  QEB_XPLAN=5,5,5 QEB_YPLAN=4,4,4 $S --env "QEB_NWP=16;QEB_NWP=24;QEB_NWP=32" >> $O/${T}_sweep.log 2>&1
echo "### all low-register" >> $O/${T}_sweep.log;         QEB_ZPLAN=8,5,5 QEB_XPLAN=5,5,5 QEB_YPLAN=4,4,4 $S --env "QEB_NWP=32,QEB_NWZ=10;QEB_NWP=32,QEB_NWZ=5" >> $O/${T}_sweep.log 2>&1
QEB_ZPLAN=8,5,5 QEB_XPLAN=5,5,5 QEB_YPLAN=4,4,4 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ausurf or h_psi" > $O/${T}_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_tests.log
tail -3 $O/${T}_tests.log; cat $O/${T}_sweep.log
