mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
rm -f gpurun_out/ch_sweep2.jsonl
for blk in 8 10 12 16; do for pol in 8,1,1 8,1,2 8,2,1 4,1,1 12,1,1; do
  echo "blocks $blk policy $pol"; RTB_TRACE_BLOCKS=$blk RTB_TRACE_POLICY=$pol timeout 300 python tools/bench_closest_hit.py --reps 2 2>&1 | grep global | sed "s/^{/{\"blocks\": $blk, \"policy\": \"$pol\", /" | tee -a gpurun_out/ch_sweep2.jsonl | cut -c1-260
done; done
for blk in 10 12 16; do
RTB_TRACE_BLOCKS=$blk timeout 600 python bench.py --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_wf_spp64_b$blk.json 2> gpurun_out/bench_wf_spp64_b$blk.err; echo "bench wf blocks $blk rc=$?"; cat gpurun_out/bench_wf_spp64_b$blk.json | cut -c1-2400
done
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 30 -c 1 -o gpurun_out/r01e_k_wf_shade $B > gpurun_out/ncu_shade.log 2>&1; tail -2 gpurun_out/ncu_shade.log
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 30 -c 1 -o gpurun_out/r01e_k_wf_trace $B > gpurun_out/ncu_trace.log 2>&1; tail -2 gpurun_out/ncu_trace.log
