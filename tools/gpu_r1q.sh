#!/bin/bash
# GPU round r1q: full parity suite, bench (ours + reference arm), launch list of the bench's own step, ncu --set full of
# the plane / z kernels at batch 15 and of the projector GEMMs at the bench's size (m = 800), other workloads
T=r1q; O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke_exit=$?" >> $O/${T}_smoke.log
python bench.py --steps 5 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench_exit=$?"
python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref_exit=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $O/${T}_ncu_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_vloc -s 3 -c 3 -f -o $O/${T}_vloc_b15 \
    python tools/prof_run.py --bands 60 --calls 1 --batch 15 > $O/${T}_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:3m' -c 2 -f -o $O/${T}_gemm_m800 \
    python tools/prof_run.py --bands 800 --calls 1 >> $O/${T}_ncu.log 2>&1
bash tools/gpu_extra.sh $T > $O/${T}_extra.log 2>&1
tail -3 $O/${T}_tests.log; tail -3 $O/${T}_smoke.log; cat $O/${T}_bench.json | cut -c1-300; cat $O/${T}_bench_ref.json | cut -c1-300; tail -2 $O/${T}_bench.err; tail -12 $O/${T}_extra.log | cut -c1-250
