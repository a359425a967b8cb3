mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/bench4_v2.log 2>&1 || true
tail -1 gpurun_out/bench4_v2.log | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2_v4.log 2>&1 || true
tail -1 gpurun_out/bench2_v4.log | cut -c1-300
