#!/bin/bash
# final validation of the committed state with the driver's own commands
T=r1z; O=gpurun_out; mkdir -p $O
python -m pytest tests -x -q -m gpu > $O/${T}_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke_exit=$?" >> $O/${T}_smoke.log
( time python bench.py --impl reference ) > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref_exit=$?"
( time python bench.py ) > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench_exit=$?"
python tools/e2e_sweep.py 0,192 > $O/${T}_e2e.log 2>&1
tail -3 $O/${T}_tests.log; tail -2 $O/${T}_smoke.log; cut -c1-250 $O/${T}_bench.json; cut -c1-250 $O/${T}_bench_ref.json; tail -4 $O/${T}_bench.err; tail -4 $O/${T}_bench_ref.err; cat $O/${T}_e2e.log
