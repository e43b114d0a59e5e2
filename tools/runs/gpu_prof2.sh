mkdir -p gpurun_out
C="python tools/bench_closest_hit.py --rays 4000000 --reps 1"
$C > gpurun_out/prof_plain_ch.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_closest_hit -s 6 -c 1 -o gpurun_out/r01b_k_closest_hit $C > gpurun_out/ncu_ch.log 2>&1
tail -3 gpurun_out/ncu_ch.log
