#!/bin/bash
# last validation of round 1: driver's commands on the final code
cd /root/repo; T=r1zr; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests rc=$?" >> $O/${T}_tests.log; tail -2 $O/${T}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${T}_smoke.log; tail -2 $O/${T}_smoke.log
timeout 600 python bench.py --impl reference > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref rc=$?"; cut -c1-160 $O/${T}_bench_ref.json
timeout 600 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"; cut -c1-200 $O/${T}_bench.json
OMP_NUM_THREADS=1 timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > $O/${T}_bench_ref_omp1.json 2>/dev/null; cut -c1-160 $O/${T}_bench_ref_omp1.json
