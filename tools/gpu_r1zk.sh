#!/bin/bash
cd /root/repo; T=r1zk; O=gpurun_out; mkdir -p $O
timeout 200 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
python -c "import json;j=json.load(open('$O/${T}_bench.json'));print(round(j['value']),round(j['e2e']['value']),j['cpu_baseline']['value'],j['roofline']['frac'],j['roofline_vloc']['frac'],j['gpu_launches'],j['clocks'])"
