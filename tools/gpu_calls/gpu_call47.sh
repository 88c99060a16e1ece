nproc; timeout 70 python tools/upload_sweep.py 256 2>&1 | tail -8
