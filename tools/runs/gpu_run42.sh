# closing single-GPU evidence for the leaf-in-link kernel at 10 CTAs/SM: tests, smoke, default bench, reference arm, launch list, ncu --set full of the trace kernel
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/r42_bench.json 2> gpurun_out/r42_bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/r42_bench.json
python bench.py --impl reference > gpurun_out/r42_bench_ref.json 2> gpurun_out/r42_bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/r42_bench_ref.json
python bench.py --workload s7_4 --no-cpu-baseline > gpurun_out/r42_bench_s7_4.json 2> gpurun_out/r42_bench_s7_4.err; echo "s7_4 rc=$?"
python tools/bench_closest_hit.py --rays 64000000 > gpurun_out/r42_closest_hit.jsonl 2> gpurun_out/r42_ch.err; echo "ch rc=$?"
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/r42_plain.json 2> gpurun_out/r42_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r01m_launches_bench.csv $B > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 14 -c 1 -o gpurun_out/r01m_k_wf_trace $B > gpurun_out/ncu_trace.log 2>&1; tail -1 gpurun_out/ncu_trace.log
