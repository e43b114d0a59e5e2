# form-3 scheduling experiments: two steps per vote (MODE 4), triangle code gated on a lane count (MODE 5); shade-kernel ncu
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r18_bench_$name.json 2> gpurun_out/r18_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r18_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), round(d["ms_per_step"],1), "trace ms", round(r["kernel_ms_total"],1), "frac", round(r["frac"],3))
PY
  env "$@" timeout 600 python tools/bench_closest_hit.py --rays 16000000 2>/dev/null | grep '"global"' > gpurun_out/r18_ch_$name.json
  python - <<PY
import json; d=json.loads(open("gpurun_out/r18_ch_$name.json").read()); print("$name soup", round(d["mrays_per_s"],1), "Mrays/s")
PY
}
run m3 RTB_TRACE_MODE=3
run m4 RTB_TRACE_MODE=4
run m5_l4 RTB_TRACE_MODE=5 RTB_TRACE_POLICY=4,1,4
run m5_l8 RTB_TRACE_MODE=5 RTB_TRACE_POLICY=4,1,8
run m5_l12 RTB_TRACE_MODE=5 RTB_TRACE_POLICY=4,1,12
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 12 -c 1 -o gpurun_out/r01j_k_wf_shade $B > gpurun_out/ncu_shade.log 2>&1; tail -1 gpurun_out/ncu_shade.log
