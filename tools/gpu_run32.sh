mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 120 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r32_bench_$name.json 2> gpurun_out/r32_bench_$name.err || echo "FAILED $name"
  python - <<PY
import json; d=json.load(open("gpurun_out/r32_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "launches", d["gpu_launches"])
PY
}
SIZE="--width 960 --height 272"
run e_s0 RTB_TRACE_SPARSE=0
run e_s10 RTB_TRACE_SPARSE=10
run e_s16_t192 RTB_TRACE_SPARSE=16 RTB_TRACE_TAIL=192
SIZE=""
run f_s10 RTB_TRACE_SPARSE=10
timeout 600 python tools/check_graph_replay.py sponza_small 640 360 24 2>&1 | tail -1
