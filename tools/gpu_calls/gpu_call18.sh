for t in 1 2 3 6; do echo "threads $t"; EIGSOL_B200_TRACE=1 EIGSOL_B200_UPLOAD_THREADS=$t python bench.py --no-cpu-baseline --steps 50 2>&1 >/dev/null | grep "staged upload"; done
echo register; EIGSOL_B200_TRACE=1 EIGSOL_B200_UPLOAD_REGISTER=1 python bench.py --no-cpu-baseline --steps 50 2>&1 >/dev/null | grep -E "staged upload|rror"
