mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2_v6.log 2>&1 || true
tail -1 gpurun_out/bench2_v6.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('device', round(d['ms_per_step'],3), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],3), round(d['e2e']['value']), d['e2e']['h2d_bytes_per_step'], d['clocks'])"
