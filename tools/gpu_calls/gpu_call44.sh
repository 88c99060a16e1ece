timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
export BLOCK_SINGLE=0
BLOCK_WIDTHS=4 timeout 200 ncu --set full --clock-control none --import-source on -k regex:gemm_block_partial -s 8 -c 1 -o gpurun_out/ncu_block_dense_k4 python tools/block_bench.py c3 3 > gpurun_out/ncu44_dense.log 2>&1; echo "dense rc=$?"
BLOCK_WIDTHS=2 timeout 200 ncu --set full --clock-control none --import-source on -k regex:spmm_csr_block -s 8 -c 1 -o gpurun_out/ncu_block_row_k2 python tools/block_bench.py c5 3 > gpurun_out/ncu44_k2.log 2>&1; echo "k2 rc=$?"
