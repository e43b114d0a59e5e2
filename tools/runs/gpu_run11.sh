mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/bench_$name.json")); r=d["roofline"]; print(round(d["value"],1), round(d["ms_per_step"],1), round(r["kernel_ms_total"],1), round(r["other_kernels_ms_total"],1), round(d["e2e"]["value"],1))
PY
}
run g1 RTB_WF_GROUPS=1
run g2 RTB_WF_GROUPS=2
run g2_b5 RTB_WF_GROUPS=2 RTB_TRACE_BLOCKS=5
run g2_b8 RTB_WF_GROUPS=2 RTB_TRACE_BLOCKS=8
