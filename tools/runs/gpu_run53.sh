# top entry of the form-3 stack in registers (RING = -1; one local load per pop instead of two) vs the shipped build
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r53_bench_$name.json 2> gpurun_out/r53_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r53_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1))
PY
  env "$@" python tools/bench_closest_hit.py --rays 16000000 2>/dev/null | grep '"global"' | cut -c1-150
}
run shipped A=1
cp raytracing-hw_b200/variants/librtb_top.so raytracing-hw_b200/librtb.so
python -m pytest tests/test_gpu_closest_hit.py tests/test_gpu_render.py -m gpu -x -q 2>&1 | tail -2
run top_b10 A=1
run top_b9 RTB_TRACE_BLOCKS=9
