cd /root/repo
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q > gpurun_out/r2p_dist_tests.log 2>&1; echo "dist rc=$?"; tail -3 gpurun_out/r2p_dist_tests.log
NG=$(python -c "import torch;print(torch.cuda.device_count())")
for F in 1 0; do
QEB_FUSED_X=$F timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $NG --mode rg --workload si512 --no-cpu-baseline --no-e2e > gpurun_out/r2p_rg_fused$F.json 2> gpurun_out/r2p_rg_fused$F.err; echo "rg fused=$F rc=$?"; cut -c1-200 gpurun_out/r2p_rg_fused$F.json
done
