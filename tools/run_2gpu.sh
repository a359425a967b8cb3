set -e
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2_v3.log 2>&1 || true
tail -1 gpurun_out/bench2_v3.log | cut -c1-300
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 bench/run_config.py --config 3 --rounds 4 > gpurun_out/cfg3_2gpu_fused_v2.log 2>&1 || true
tail -1 gpurun_out/cfg3_2gpu_fused_v2.log | cut -c1-600
