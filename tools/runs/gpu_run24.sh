# round-1 final single-GPU evidence: full default bench, reference arm, launch list, ncu --set full of the trace and shade kernels
mkdir -p gpurun_out
python bench.py > gpurun_out/r24_bench.json 2> gpurun_out/r24_bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/r24_bench.json
python bench.py --impl reference > gpurun_out/r24_bench_ref.json 2> gpurun_out/r24_bench_ref.err; echo "ref rc=$?"
python bench.py --workload s7_4 --no-cpu-baseline > gpurun_out/r24_bench_s7_4.json 2> gpurun_out/r24_bench_s7_4.err; echo "s7_4 rc=$?"
python bench.py --workload soup_10m --spp 8 --no-cpu-baseline > gpurun_out/r24_bench_soup10m.json 2> gpurun_out/r24_bench_soup10m.err; echo "soup10m rc=$?"
python tools/bench_closest_hit.py --rays 64000000 > gpurun_out/r24_closest_hit.jsonl 2> gpurun_out/r24_ch.err; echo "ch rc=$?"
python tools/bench_closest_hit.py --tris 10000000 --half 21.5 --rays 16000000 >> gpurun_out/r24_closest_hit.jsonl 2>> gpurun_out/r24_ch.err
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/r24_plain.json 2> gpurun_out/r24_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r01k_launches_bench.csv $B > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 14 -c 1 -o gpurun_out/r01k_k_wf_trace $B > gpurun_out/ncu_trace.log 2>&1; tail -1 gpurun_out/ncu_trace.log
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 14 -c 1 -o gpurun_out/r01k_k_wf_shade $B > gpurun_out/ncu_shade.log 2>&1; tail -1 gpurun_out/ncu_shade.log
