# suspension of tail rays (RTB_TRACE_TAIL): parity, equality with the unsuspended frame, sweep at 1/8 canvas and full canvas
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r22_pytest.log; cat gpurun_out/r22_pytest.log
timeout 900 python tools/check_graph_replay.py sponza 640 360 24 2>&1 | cut -c1-60,120-200 | tee gpurun_out/r22_check.log
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r22_bench_$name.json 2> gpurun_out/r22_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r22_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace ms", round(r["kernel_ms_total"],1), "launches", d["gpu_launches"])
PY
}
SIZE="--width 960 --height 272"
run small_t0 RTB_TRACE_TAIL=0
run small_t48 RTB_TRACE_TAIL=48
run small_t96 RTB_TRACE_TAIL=96
run small_t160 RTB_TRACE_TAIL=160
run small_t96_g4 RTB_TRACE_TAIL=96 RTB_WF_GROUPS=4
run small_t96_g2 RTB_TRACE_TAIL=96 RTB_WF_GROUPS=2
SIZE=""
run full_t0 RTB_TRACE_TAIL=0
run full_t96 RTB_TRACE_TAIL=96
run full_t160 RTB_TRACE_TAIL=160
