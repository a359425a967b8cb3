mkdir -p gpurun_out
for i in 1 2 3; do timeout 300 python -m pytest tests/test_b200_learner_gpu.py -m gpu -q 2>&1 | grep -E "^E  |passed|failed" | head -8; done
