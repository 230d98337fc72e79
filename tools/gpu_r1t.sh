#!/bin/bash
# 8-GPU box: distributed parity (4 ranks) incl. p2p, Si512 R&G x8 / x4 with both exchanges
T=r1t; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -k multi_gpu > $O/${T}_dist_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_dist_tests.log
tail -5 $O/${T}_dist_tests.log
for n in 8 4; do for ex in nccl p2p; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus $n --steps 5 --warmup 3 --mode rg --workload si512 --exchange $ex > $O/${T}_n${n}_rg_$ex.json 2> $O/${T}_n${n}_rg_$ex.err
  echo "n=$n $ex exit=$?"; python -c "import json;j=json.load(open('$O/${T}_n${n}_rg_$ex.json'));print(j['value'],j['ms_per_step'],j['sections_ms'])"; tail -2 $O/${T}_n${n}_rg_$ex.err
done; done
