set -e
mkdir -p gpurun_out
timeout 200 python tools/profile_target.py mlpstep 2>&1 | grep -A8 "cluster=16" | head -40
timeout 300 python -m pytest tests/test_b200_learner_gpu.py -x -q -k "chain_kernel or pdl or mlp" 2>&1 | tail -3
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench1_v8.log 2>&1 || true
tail -1 gpurun_out/bench1_v8.log | cut -c1-250
