#!/bin/bash
# noncollinear h_psi: parity suite
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1zw_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r1zw_tests.log
tail -25 gpurun_out/r1zw_tests.log
