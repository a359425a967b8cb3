set -e
mkdir -p gpurun_out
echo "--- n_iter=1"; timeout 200 python tools/profile_target.py mlpstep 2>&1 | grep -A7 "cluster=16" | tail -7
echo "--- n_iter=4 (stamps of the last pass)"; P2PFL_CHAIN_ITER=4 timeout 200 python tools/profile_target.py mlpstep 2>&1 | grep -A7 "cluster=16" | tail -7
