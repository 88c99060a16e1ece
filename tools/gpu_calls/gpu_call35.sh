export BLOCK_WIDTHS=4 BLOCK_SINGLE=0
BLOCK_CONFIGS=tile timeout 240 ncu --set full --clock-control none --import-source on -k regex:spmm_csr_tile -s 8 -c 1 -o gpurun_out/ncu_block_tile_k4 python tools/block_bench.py c5 3 > gpurun_out/ncu35_tile.log 2>&1; echo "tile rc=$?"
BLOCK_CONFIGS=row timeout 240 ncu --set full --clock-control none --import-source on -k regex:spmm_csr_block -s 8 -c 1 -o gpurun_out/ncu_block_row_k4 python tools/block_bench.py c5 3 > gpurun_out/ncu35_row.log 2>&1; echo "row rc=$?"
ls -la gpurun_out/*.ncu-rep
