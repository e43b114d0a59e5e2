mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r20_bench_$name.json 2> gpurun_out/r20_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r20_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; launches", d["gpu_launches"])
PY
}
SIZE="--width 960 --height 272"
run small_g8 RTB_WF_GROUPS=8
run small_g12 RTB_WF_GROUPS=12
run small_g16 RTB_WF_GROUPS=16
run small_g16_b6 RTB_WF_GROUPS=16 RTB_TRACE_BLOCKS=6
SIZE=""
run full_g2 RTB_WF_GROUPS=2
run full_g4 RTB_WF_GROUPS=4
