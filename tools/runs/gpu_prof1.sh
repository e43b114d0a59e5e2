mkdir -p gpurun_out
set -x
B="python bench.py --spp 4 --steps 2 --warmup 1 --no-cpu-baseline"
C="python tools/bench_closest_hit.py --rays 4000000 --reps 1"
$B > gpurun_out/prof_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_bench.csv $B > gpurun_out/ncu_launches.log 2>&1
$B > gpurun_out/prof_plain_bench2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render -s 1 -c 1 -o gpurun_out/r01_k_render $B > gpurun_out/ncu_render.log 2>&1
$C > gpurun_out/prof_plain_ch.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_closest_hit -s 2 -c 1 -o gpurun_out/r01_k_closest_hit $C > gpurun_out/ncu_ch.log 2>&1
ls -la gpurun_out
