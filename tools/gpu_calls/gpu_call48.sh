timeout 28 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 25 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "ingest or reference_" 2>&1 | tail -2
