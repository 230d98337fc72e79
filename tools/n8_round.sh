#!/bin/bash
# 8-GPU validation of one round: tools/n8_round.sh TAG  (dist parity tests on all GPUs, the default bench line at N = all GPUs
# with its strong-scaling modes and Davidson solve, the R&G Si512 line with the fused and the unfused exchange)
T=${1:-rX}; O=gpurun_out; mkdir -p $O; cd "$(dirname "$0")/.."
NG=$(python -c "import torch;print(torch.cuda.device_count())")
timeout 1200 python -m pytest tests/test_gpu_dist.py -m gpu -x -q > $O/${T}_n${NG}_dist_tests.log 2>&1; echo "dist rc=$?"; tail -3 $O/${T}_n${NG}_dist_tests.log
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $NG"
timeout 900 $RUN --no-cpu-baseline > $O/${T}_n${NG}.json 2> $O/${T}_n${NG}.err; echo "bench N=$NG rc=$?"; cut -c1-200 $O/${T}_n${NG}.json
for F in 1 0; do
  QEB_FUSED_X=$F timeout 600 $RUN --mode rg --workload si512 --no-cpu-baseline --no-e2e --no-davidson --no-strong > $O/${T}_n${NG}_rg_fused$F.json 2> $O/${T}_n${NG}_rg_fused$F.err
  echo "rg fused=$F rc=$?"; cut -c1-200 $O/${T}_n${NG}_rg_fused$F.json
done
