# top of the form-3 stack in a shared-memory ring (RTB_RING entries per thread): parity with the default build, then ring 0 / 4 / 8 / 16
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
cp raytracing-hw_b200/librtb.so raytracing-hw_b200/variants/librtb_r8.so
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r43_bench_$name.json 2> gpurun_out/r43_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r43_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1))
PY
  env "$@" python tools/bench_closest_hit.py --rays 16000000 2>/dev/null | grep '"global"' | cut -c1-150
}
for r in 0 4 8 16; do cp raytracing-hw_b200/variants/librtb_r$r.so raytracing-hw_b200/librtb.so; run ring${r}_b10 RTB_TRACE_BLOCKS=10; done
for r in 8 16; do cp raytracing-hw_b200/variants/librtb_r$r.so raytracing-hw_b200/librtb.so; run ring${r}_b9 RTB_TRACE_BLOCKS=9; done
