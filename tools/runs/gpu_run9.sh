mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_closest_hit.py --reps 2 2>&1 | tee gpurun_out/ch_1m_16m.jsonl | cut -c1-250
timeout 600 python tools/bench_closest_hit.py --rays 64000000 --reps 2 2>&1 | tee gpurun_out/ch_1m_64m.jsonl | cut -c1-250
timeout 900 python tools/bench_closest_hit.py --tris 10000000 --half 21.5 --rays 16000000 --reps 2 2>&1 | tee gpurun_out/ch_10m_16m.jsonl | cut -c1-250
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; cat gpurun_out/bench_full.json; tail -3 gpurun_out/bench_full.err
timeout 600 python bench.py --workload s7_4 --no-cpu-baseline > gpurun_out/bench_s7_4.json 2> gpurun_out/bench_s7_4.err; echo "bench s7_4 rc=$?"; cat gpurun_out/bench_s7_4.json | cut -c1-1500
timeout 900 python bench.py --workload soup_10m --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_soup10m_spp8.json 2> gpurun_out/bench_soup10m.err; echo "bench soup10m rc=$?"; cat gpurun_out/bench_soup10m_spp8.json | cut -c1-2500; tail -3 gpurun_out/bench_soup10m.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; cat gpurun_out/bench_reference.json
