# CUDA-graph replay of the groups' iterations: equality with direct launches, then the 1/8-canvas sweep
mkdir -p gpurun_out
timeout 900 python tools/check_graph_replay.py sponza 640 360 24 2>&1 | tee gpurun_out/r21_check_graph.log
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r21_bench_$name.json 2> gpurun_out/r21_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r21_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; e2e", round(d["e2e"]["value"],1), "launches", d["gpu_launches"])
PY
}
SIZE="--width 960 --height 272"
run small_g8_direct RTB_WF_GROUPS=8 RTB_WF_GRAPH=0
run small_g8 RTB_WF_GROUPS=8
run small_g12 RTB_WF_GROUPS=12
run small_g16 RTB_WF_GROUPS=16
run small_g4 RTB_WF_GROUPS=4
SIZE=""
run full_g2 RTB_WF_GROUPS=2
