mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/bench_closest_hit.py > gpurun_out/ch_1m.json 2>&1; cat gpurun_out/ch_1m.json
timeout 600 python bench.py --spp 16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_spp16.json 2> gpurun_out/bench_spp16.err; echo "bench rc=$?"; cat gpurun_out/bench_spp16.json; tail -5 gpurun_out/bench_spp16.err
