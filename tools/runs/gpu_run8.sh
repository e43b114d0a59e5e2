mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
rm -f gpurun_out/ch_sweep4.jsonl
for mode in 1 2; do for blk in 9 10 12; do for pol in 4,1,1 4,1,4 4,1,8; do
  echo "mode $mode blocks $blk policy $pol"; RTB_TRACE_MODE=$mode RTB_TRACE_BLOCKS=$blk RTB_TRACE_POLICY=$pol timeout 300 python tools/bench_closest_hit.py --reps 2 2>&1 | grep global | sed "s/^{/{\"mode\": $mode, \"blocks\": $blk, \"policy\": \"$pol\", /" | tee -a gpurun_out/ch_sweep4.jsonl | cut -c1-200
done; done; done
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/bench_$name.json")); r=d["roofline"]; print(round(d["value"],1), round(d["ms_per_step"],1), round(r["kernel_ms_total"],1), round(r["other_kernels_ms_total"],1), round(r["kernel_mrays_per_s"],1))
PY
}
run m1_b10_s5 RTB_TRACE_MODE=1 RTB_TRACE_BLOCKS=10 RTB_TRACE_POLICY=4,1,1 RTB_SHADE_BLOCKS=5
run m1_b10_s6 RTB_TRACE_MODE=1 RTB_TRACE_BLOCKS=10 RTB_TRACE_POLICY=4,1,1 RTB_SHADE_BLOCKS=6
run m1_b10_s8 RTB_TRACE_MODE=1 RTB_TRACE_BLOCKS=10 RTB_TRACE_POLICY=4,1,1 RTB_SHADE_BLOCKS=8
run m2_b10_s6 RTB_TRACE_MODE=2 RTB_TRACE_BLOCKS=10 RTB_TRACE_POLICY=4,1,1 RTB_SHADE_BLOCKS=6
run m1_b9_s6 RTB_TRACE_MODE=1 RTB_TRACE_BLOCKS=9 RTB_TRACE_POLICY=4,1,1 RTB_SHADE_BLOCKS=6
run m1_b10_l4_s6 RTB_TRACE_MODE=1 RTB_TRACE_BLOCKS=10 RTB_TRACE_POLICY=4,1,4 RTB_SHADE_BLOCKS=6
