set -e
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "mn_major or split_k or conv_backward_pair or gemm_tn_plain or grouped" 2>&1 | tail -15
timeout 300 python -m pytest tests/test_conv_kernels_gpu.py -x -q 2>&1 | tail -15
timeout 300 python -m pytest tests/test_b200_learner_gpu.py -x -q -k resnet -s 2>&1 | tail -8
timeout 200 python tools/profile_target.py resnet > gpurun_out/resnet_plain.log 2>&1 || true
tail -2 gpurun_out/resnet_plain.log
