set -e
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_conv_kernels_gpu.py -x -q -k "gemm or conv" 2>&1 | tail -5
timeout 200 python tools/profile_target.py resnet > gpurun_out/resnet_plain.log 2>&1
tail -1 gpurun_out/resnet_plain.log
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/resnet_launches_v3.csv python tools/profile_target.py resnet > gpurun_out/resnet_ncu.log 2>&1 || true
tail -1 gpurun_out/resnet_ncu.log
