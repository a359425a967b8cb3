mkdir -p gpurun_out
timeout 130 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29661 bench/run_config.py --config 4 --rounds 3 --timeout 100 > gpurun_out/cfg4_8gpu_64.log 2>&1 || true
tail -2 gpurun_out/cfg4_8gpu_64.log | cut -c1-600
