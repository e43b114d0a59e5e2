mkdir -p gpurun_out
nvidia-smi -L | head -2; nproc
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/bench_closest_hit.py > gpurun_out/ch_1m.json 2>&1; cat gpurun_out/ch_1m.json
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; cat gpurun_out/bench_full.json; tail -5 gpurun_out/bench_full.err
C="python tools/bench_closest_hit.py --rays 4000000 --reps 1"
ncu --set full --clock-control none --import-source on -k regex:k_closest_hit_queue -s 2 -c 1 -o gpurun_out/r01c_k_closest_hit_queue $C > gpurun_out/ncu_ch.log 2>&1
tail -3 gpurun_out/ncu_ch.log
