#!/bin/bash
# 8-GPU box, final code: k-pool bench with the driver's command, Si512 R&G p2p
cd /root/repo; T=r1zo; O=gpurun_out; mkdir -p $O
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus 8 --steps 5 --warmup 3 > $O/${T}_n8_pools.json 2> $O/${T}_n8_pools.err; echo "pools exit=$?"
cut -c1-220 $O/${T}_n8_pools.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 \
    bench.py --gpus 8 --steps 3 --warmup 3 --mode rg --workload si512 > $O/${T}_n8_rg_p2p.json 2> $O/${T}_n8_rg_p2p.err; echo "rg exit=$?"
cut -c1-220 $O/${T}_n8_rg_p2p.json
