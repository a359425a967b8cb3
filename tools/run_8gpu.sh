mkdir -p gpurun_out
export P2PFL_BENCH_TRACE=gpurun_out/trace8b
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench8_v5.log 2>&1 || true
tail -1 gpurun_out/bench8_v5.log | cut -c1-300
