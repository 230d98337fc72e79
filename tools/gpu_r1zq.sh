#!/bin/bash
cd /root/repo; T=r1zq; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests rc=$?" >> $O/${T}_tests.log; tail -3 $O/${T}_tests.log
for sub in 0 40 20 80; do
  QEB_PIPE_SUB=$sub timeout 600 python bench.py --no-cpu-baseline > $O/${T}_bench_sub$sub.json 2> /dev/null
  python -c "import json;j=json.load(open('$O/${T}_bench_sub$sub.json'));print('sub',$sub,'value',round(j['value']),'e2e',round(j['e2e']['value']),j['e2e']['ms_per_step'])"
done
