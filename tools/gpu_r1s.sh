#!/bin/bash
# 2-GPU box: distributed parity incl. the peer-to-peer exchange, then Si512 R&G with both exchanges
T=r1s; O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -k multi_gpu > $O/${T}_dist_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_dist_tests.log
tail -25 $O/${T}_dist_tests.log
for ex in nccl p2p; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus 2 --steps 3 --warmup 3 --mode rg --workload si512 --exchange $ex > $O/${T}_n2_rg_$ex.json 2> $O/${T}_n2_rg_$ex.err
  echo "$ex exit=$?"; cut -c1-200 $O/${T}_n2_rg_$ex.json; python -c "import json;j=json.load(open('$O/${T}_n2_rg_$ex.json'));print(j['value'],j['ms_per_step'],j['sections_ms'])"; tail -3 $O/${T}_n2_rg_$ex.err
done
