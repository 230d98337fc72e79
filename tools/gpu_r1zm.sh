#!/bin/bash
cd /root/repo; T=r1zm; O=gpurun_out; mkdir -p $O
for ch in 0 256 272 400; do
  QEB_PIPELINE_CHUNK=$ch timeout 300 python bench.py --no-cpu-baseline --steps 4 > $O/${T}_bench_chunk$ch.json 2> /dev/null
  python -c "import json;j=json.load(open('$O/${T}_bench_chunk$ch.json'));print('chunk',$ch,'value',round(j['value']),'e2e',round(j['e2e']['value']),j['e2e']['ms_per_step'])"
done
