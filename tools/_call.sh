python tools/gpu_shard_stats.py 10k 8 2>&1 | tail -1
python tools/gpu_stats.py 10k 2>&1 | grep render | cut -c1-160
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/s16k_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s16k_pytest_gpu.log
timeout 300 python bench.py > gpurun_out/s16k_bench.json 2> gpurun_out/s16k_bench.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/s16k_bench.json
