# closing single-GPU evidence on the final code: smoke, default bench (with cpu baseline), reference arm
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/r52_bench.json 2> gpurun_out/r52_bench.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/r52_bench.json
python bench.py --impl reference > gpurun_out/r52_bench_ref.json 2> gpurun_out/r52_bench_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/r52_bench_ref.json
