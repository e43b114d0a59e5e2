# small-canvas sweep of rays per lane x groups with tail suspension on
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r23_bench_$name.json 2> gpurun_out/r23_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r23_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace ms", round(r["kernel_ms_total"],1), "launches", d["gpu_launches"])
PY
}
SIZE="--width 960 --height 272"
run g8_rpl2 RTB_TRACE_RPL=2
run g8_rpl3 RTB_TRACE_RPL=3
run g4_rpl2 RTB_TRACE_RPL=2 RTB_WF_GROUPS=4
run g4_rpl3 RTB_TRACE_RPL=3 RTB_WF_GROUPS=4
run g2_rpl4 RTB_TRACE_RPL=4 RTB_WF_GROUPS=2
run g8_rpl2_t64 RTB_TRACE_RPL=2 RTB_TRACE_TAIL=64
run g8_rpl2_t32 RTB_TRACE_RPL=2 RTB_TRACE_TAIL=32
SIZE=""
run full_rpl2 RTB_TRACE_RPL=2
run full_t96_b10 RTB_TRACE_BLOCKS=10
