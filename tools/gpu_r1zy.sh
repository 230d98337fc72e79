#!/bin/bash
# spinor x pass with the static plans: parity suite + timing at AUSURF112 size
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1zy_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r1zy_tests.log
tail -3 gpurun_out/r1zy_tests.log
timeout 600 python tools/bench_extras.py > gpurun_out/r1zy_extras.jsonl 2> gpurun_out/r1zy_extras.err; echo "extras rc=$?"
tail -1 gpurun_out/r1zy_extras.jsonl
QEB_NWX=8 timeout 600 python tools/bench_extras.py 2>/dev/null | tail -1
QEB_NWX=16 timeout 600 python tools/bench_extras.py 2>/dev/null | tail -1
