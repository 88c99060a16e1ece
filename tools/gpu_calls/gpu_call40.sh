export BLOCK_SINGLE=0
BLOCK_WIDTHS=4 timeout 240 ncu --set full --clock-control none --import-source on -k regex:spmm_csr_tile -s 8 -c 1 -o gpurun_out/ncu_block_tile2_k4 python tools/block_bench.py c5 3 > gpurun_out/ncu40_k4.log 2>&1; echo "k4 rc=$?"
BLOCK_WIDTHS=8 timeout 240 ncu --set full --clock-control none --import-source on -k regex:spmm_csr_tile -s 8 -c 1 -o gpurun_out/ncu_block_tile2_k8 python tools/block_bench.py c5 3 > gpurun_out/ncu40_k8.log 2>&1; echo "k8 rc=$?"
