# consolidation of the tail's rays through the in-launch requeue: parity, equality across variants, sweep
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python tools/check_graph_replay.py sponza 640 360 24 2>&1 | cut -c1-40,100-200 | tee gpurun_out/r31_check.log
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r31_bench_$name.json 2> gpurun_out/r31_bench_$name.err || echo "FAILED $name"
  python - <<PY
import json; d=json.load(open("gpurun_out/r31_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "launches", d["gpu_launches"])
PY
}
SIZE="--width 960 --height 272"
run e_s0 RTB_TRACE_SPARSE=0
run e_s6 RTB_TRACE_SPARSE=6
run e_s10 RTB_TRACE_SPARSE=10
run e_s16 RTB_TRACE_SPARSE=16
run e_s10_t192 RTB_TRACE_SPARSE=10 RTB_TRACE_TAIL=192
run e_s16_t192 RTB_TRACE_SPARSE=16 RTB_TRACE_TAIL=192
run e_s16_t512 RTB_TRACE_SPARSE=16 RTB_TRACE_TAIL=512
run e_s16_t192_g4 RTB_TRACE_SPARSE=16 RTB_TRACE_TAIL=192 RTB_WF_GROUPS=4
SIZE=""
run f_s0 RTB_TRACE_SPARSE=0
run f_s10 RTB_TRACE_SPARSE=10
run f_s16_t192 RTB_TRACE_SPARSE=16 RTB_TRACE_TAIL=192
