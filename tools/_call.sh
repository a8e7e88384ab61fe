python tools/gpu_stats.py 10k 2>&1 | grep -v "^heavy"
python tools/gpu_stats.py 1m 2>&1 | grep -v "^heavy"
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
bash tools/gpu_ncu_src.sh s16f k_shade 10
