timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu26.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/pytest_gpu26.log | cut -c1-400
