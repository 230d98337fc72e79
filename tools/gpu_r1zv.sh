#!/bin/bash
# 2-GPU box, final code: distributed parity suite, k-pool bench (the driver's N=2 command), Si512 R&G p2p
T=r1zv; O=gpurun_out; mkdir -p $O
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -k multi_gpu > $O/${T}_dist_tests.log 2>&1; echo "tests_exit=$?" >> $O/${T}_dist_tests.log
tail -4 $O/${T}_dist_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus 2 --steps 5 --warmup 3 > $O/${T}_n2_pools.json 2> $O/${T}_n2_pools.err; echo "pools exit=$?"
cut -c1-250 $O/${T}_n2_pools.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 \
    bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/${T}_n2_ref.json 2> $O/${T}_n2_ref.err; echo "ref exit=$?"
cut -c1-200 $O/${T}_n2_ref.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29515 \
    bench.py --gpus 2 --steps 3 --warmup 3 --mode rg --workload si512 > $O/${T}_n2_rg_p2p.json 2> $O/${T}_n2_rg_p2p.err; echo "rg exit=$?"
cut -c1-250 $O/${T}_n2_rg_p2p.json
