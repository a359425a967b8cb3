set -e
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_b200_learner_gpu.py -x -q -k "stream_mode or fit_and_evaluate" 2>&1 | tail -8
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench1_v12.log 2>&1 || true
tail -1 gpurun_out/bench1_v12.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('device', round(d['ms_per_step'],3), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],3), round(d['e2e']['value']), d['e2e']['h2d_bytes_per_step'])"
