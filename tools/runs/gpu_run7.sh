mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
rm -f gpurun_out/ch_sweep3.jsonl
for mode in 0 1; do for blk in 8 10 12 16; do for pol in 8,1,1 4,1,1 16,1,1; do
  if [ $mode = 0 ] && [ $pol != 8,1,1 ]; then continue; fi
  echo "mode $mode blocks $blk policy $pol"; RTB_TRACE_MODE=$mode RTB_TRACE_BLOCKS=$blk RTB_TRACE_POLICY=$pol timeout 300 python tools/bench_closest_hit.py --reps 2 2>&1 | grep global | sed "s/^{/{\"mode\": $mode, \"blocks\": $blk, \"policy\": \"$pol\", /" | tee -a gpurun_out/ch_sweep3.jsonl | cut -c1-230
done; done; done
for mode in 0 1; do for blk in 10 12; do
RTB_TRACE_MODE=$mode RTB_TRACE_BLOCKS=$blk timeout 600 python bench.py --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_wf_spp64_m${mode}_b$blk.json 2> gpurun_out/bench_wf_spp64_m${mode}_b$blk.err; echo "bench wf mode $mode blocks $blk rc=$?"; python - <<PY
import json; d=json.load(open("gpurun_out/bench_wf_spp64_m${mode}_b$blk.json")); r=d["roofline"]; print(d["value"], d["ms_per_step"], r["kernel_ms_total"], r["other_kernels_ms_total"], r["kernel_mrays_per_s"], r["box_tests_per_ray"], r["tri_tests_per_ray"])
PY
done; done
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 12 -c 1 -o gpurun_out/r01f_k_wf_trace $B > gpurun_out/ncu_trace.log 2>&1; tail -2 gpurun_out/ncu_trace.log
