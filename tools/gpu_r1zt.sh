#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "nc" > gpurun_out/r1zs_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r1zs_tests.log
tail -3 gpurun_out/r1zs_tests.log
timeout 600 python tools/bench_extras.py 2>/dev/null | tail -1 | tee gpurun_out/r1zs_nc.json
QEB_XSPINX=0 timeout 600 python tools/bench_extras.py 2>/dev/null | tail -1
QEB_NWS=16 timeout 600 python tools/bench_extras.py 2>/dev/null | tail -1
QEB_NWS=8 timeout 600 python tools/bench_extras.py 2>/dev/null | tail -1
