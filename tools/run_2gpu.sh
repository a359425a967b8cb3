mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu15_2gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu15_2gpu.log
tail -3 gpurun_out/pytest_gpu15_2gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2_v5.log 2>&1 || true
tail -1 gpurun_out/bench2_v5.log | cut -c1-260
