mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/bench_$name.json")); r=d["roofline"]; print(round(d["value"],1), round(d["ms_per_step"],1), round(d["e2e"]["value"],1))
PY
}
SIZE="--width 960 --height 272"
run small_g1 RTB_WF_GROUPS=1
run small_g2 RTB_WF_GROUPS=2
run small_g4 RTB_WF_GROUPS=4
run small_g8 RTB_WF_GROUPS=8
run small_g4_b5 RTB_WF_GROUPS=4 RTB_TRACE_BLOCKS=5
SIZE=""
run full_g4 RTB_WF_GROUPS=4
run full_g8 RTB_WF_GROUPS=8
