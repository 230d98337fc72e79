#!/bin/bash
# GPU round r1o: register-budget variants of the plane kernel (320 threads / 10 warps) and of the pass kernels
T=r1o; O=gpurun_out; mkdir -p $O
echo "### default lib" >> $O/${T}_sweep.log
python tools/vloc_sweep.py --bands 240 --batches 15 --env "QEB_NWP=16;QEB_NWP=10" >> $O/${T}_sweep.log 2>&1
echo "### p320 (plane kernel __launch_bounds__(320))" >> $O/${T}_sweep.log
QEB_LIB=q-e_b200/libqeb200_p320.so python tools/vloc_sweep.py --bands 240 --batches 15 --env "QEB_NWP=10;QEB_NWP=8;QEB_NWP=10,QEB_PLANE_BPC=5" >> $O/${T}_sweep.log 2>&1
echo "### f256 (pass kernels __launch_bounds__(256), plane 320)" >> $O/${T}_sweep.log
QEB_LIB=q-e_b200/libqeb200_f256.so python tools/vloc_sweep.py --bands 240 --batches 15 --env "QEB_NWP=10;QEB_NWP=10,QEB_NWZ=5;QEB_NWP=10,QEB_NWZ=8;QEB_PLANE=0" >> $O/${T}_sweep.log 2>&1
cat $O/${T}_sweep.log
