# 1/8-canvas emulation (what one of 8 GPUs renders of a 1080p frame) vs the full canvas, current defaults; group sweep
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r19_bench_$name.json 2> gpurun_out/r19_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r19_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace ms", round(r["kernel_ms_total"],1), "launches", d["gpu_launches"])
PY
}
SIZE=""
run full
SIZE="--width 960 --height 272"
run small_auto
run small_g1 RTB_WF_GROUPS=1
run small_g4 RTB_WF_GROUPS=4
run small_g8 RTB_WF_GROUPS=8
run small_g8_b5 RTB_WF_GROUPS=8 RTB_TRACE_BLOCKS=5
