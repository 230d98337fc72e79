#!/bin/bash
# final validation of round 1 with the driver's commands + the noncollinear V_loc test and timing
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1zz_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r1zz_tests.log
tail -3 gpurun_out/r1zz_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1zz_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r1zz_smoke.log
tail -2 gpurun_out/r1zz_smoke.log
timeout 600 python bench.py --impl reference > gpurun_out/r1zz_bench_ref.json 2> gpurun_out/r1zz_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py > gpurun_out/r1zz_bench.json 2> gpurun_out/r1zz_bench.err; echo "bench rc=$?"
cat gpurun_out/r1zz_bench.json | cut -c1-400
timeout 600 python tools/bench_extras.py > gpurun_out/r1zz_extras.jsonl 2> gpurun_out/r1zz_extras.err; echo "extras rc=$?"
tail -2 gpurun_out/r1zz_extras.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1zz_launches.csv python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r1zz_ncu.log 2>&1; echo "ncu rc=$?"
