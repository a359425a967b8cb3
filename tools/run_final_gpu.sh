mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu14.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu14.log
tail -3 gpurun_out/pytest_gpu14.log
timeout 200 python -c "import __graft_entry__ as e; e.smoke(); print('smoke ok')" 2>&1 | tail -2
