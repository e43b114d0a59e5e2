mkdir -p gpurun_out
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/prof_plain_bench.json 2> gpurun_out/prof_plain_bench.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r01h_launches_bench.csv $B > gpurun_out/ncu_launches.log 2>&1
cat gpurun_out/prof_plain_bench.json | cut -c1-300
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 12 -c 1 -o gpurun_out/r01h_k_wf_trace $B > gpurun_out/ncu_trace.log 2>&1; tail -1 gpurun_out/ncu_trace.log
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 12 -c 1 -o gpurun_out/r01h_k_wf_shade $B > gpurun_out/ncu_shade.log 2>&1; tail -1 gpurun_out/ncu_shade.log
