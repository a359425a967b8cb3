set -e
mkdir -p gpurun_out
export P2PFL_BENCH_TRACE=gpurun_out/trace8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench8_v4.log 2>&1 || true
tail -1 gpurun_out/bench8_v4.log | cut -c1-300
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench/run_config.py --config 3 --rounds 4 > gpurun_out/cfg3_8gpu_fused.log 2>&1 || true
tail -1 gpurun_out/cfg3_8gpu_fused.log | cut -c1-600
nproc > gpurun_out/nproc.txt
