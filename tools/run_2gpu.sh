mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641 bench/run_config.py --config 4 --rounds 4 --timeout 250 > gpurun_out/cfg4_2gpu_16_v2.log 2>&1 || true
tail -1 gpurun_out/cfg4_2gpu_16_v2.log | cut -c1-500
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29651 bench/run_config.py --config 4 --per-gpu 16 --rounds 3 --timeout 250 > gpurun_out/cfg4_2gpu_32.log 2>&1 || true
tail -1 gpurun_out/cfg4_2gpu_32.log | cut -c1-500
