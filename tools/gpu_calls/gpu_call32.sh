timeout 200 python -m pytest tests/test_gpu_block.py -m gpu -q -x --timeout 120 > gpurun_out/pytest_block32.log 2>&1; echo "block tests rc=$?"; tail -25 gpurun_out/pytest_block32.log
timeout 170 python tools/block_bench.py c5 40 > gpurun_out/block_bench32_c5.jsonl 2> gpurun_out/block_bench32_c5.err; echo "bench rc=$?"; cat gpurun_out/block_bench32_c5.jsonl; tail -3 gpurun_out/block_bench32_c5.err
