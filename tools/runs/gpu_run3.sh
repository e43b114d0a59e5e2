mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_closest_hit.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_closest_hit.py > gpurun_out/ch_1m.json 2>&1; cat gpurun_out/ch_1m.json
