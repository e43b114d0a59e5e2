# groups x canvas share sweep (1/2, 1/4, 1/8 of a 1080p frame on one GPU), tail suspension on
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r27_bench_$name.json 2> gpurun_out/r27_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r27_bench_$name.json")); print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; launches", d["gpu_launches"])
PY
}
SIZE="--width 1920 --height 544"
for g in 1 2 4; do run half_g$g RTB_WF_GROUPS=$g; done
SIZE="--width 960 --height 544"
for g in 1 2 4 8; do run quarter_g$g RTB_WF_GROUPS=$g; done
run quarter_g2_t160 RTB_WF_GROUPS=2 RTB_TRACE_TAIL=160
run quarter_g4_t160 RTB_WF_GROUPS=4 RTB_TRACE_TAIL=160
SIZE="--width 960 --height 272"
for g in 2 4 6 8; do run eighth_g$g RTB_WF_GROUPS=$g; done
run eighth_g6_t128 RTB_WF_GROUPS=6 RTB_TRACE_TAIL=128
