#!/bin/bash
cd /root/repo; T=r1zp; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests rc=$?" >> $O/${T}_tests.log; tail -3 $O/${T}_tests.log
for tail in 0 64 128; do
  QEB_PIPE_TAIL=$tail timeout 600 python bench.py --no-cpu-baseline > $O/${T}_bench_tail$tail.json 2> /dev/null
  python -c "import json;j=json.load(open('$O/${T}_bench_tail$tail.json'));print('tail',$tail,'value',round(j['value']),'e2e',round(j['e2e']['value']),j['e2e']['ms_per_step'])"
done
