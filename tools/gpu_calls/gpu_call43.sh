timeout 400 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu43.log 2>&1; echo "gpu suite rc=$?"; tail -4 gpurun_out/pytest_gpu43.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python bench.py > gpurun_out/bench43_c5.json 2> gpurun_out/bench43_c5.err; echo "bench rc=$?"; cat gpurun_out/bench43_c5.json | cut -c1-900
BLOCK_WIDTHS=1,2,4,8 timeout 150 python tools/block_bench.py c3 30 > gpurun_out/block_bench43_c3.jsonl 2> gpurun_out/block_bench43_c3.err; echo "c3 rc=$?"; cut -c1-140 gpurun_out/block_bench43_c3.jsonl
BLOCK_WIDTHS=1,2,4 timeout 150 python tools/block_bench.py c4 30 > gpurun_out/block_bench43_c4.jsonl 2> gpurun_out/block_bench43_c4.err; echo "c4 rc=$?"; cut -c1-140 gpurun_out/block_bench43_c4.jsonl
