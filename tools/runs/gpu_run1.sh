mkdir -p gpurun_out
set -x
nvidia-smi -L | head -2; nproc; free -g | head -2
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --spp 16 --steps 2 --warmup 3 > gpurun_out/bench_spp16.json 2> gpurun_out/bench_spp16.err; echo "bench rc=$?"; cat gpurun_out/bench_spp16.json; tail -5 gpurun_out/bench_spp16.err
timeout 600 python tools/bench_closest_hit.py > gpurun_out/ch_1m.json 2>&1; cat gpurun_out/ch_1m.json
