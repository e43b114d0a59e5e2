# progressive rendering (rtb_render_resume, progress callback, CLI knobs): the whole GPU suite, then the default bench for regressions
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r46_bench.json 2> gpurun_out/r46_bench.err; echo "bench rc=$?"
python - <<PY
import json; d=json.load(open("gpurun_out/r46_bench.json")); r=d["roofline"]; print(round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1), "e2e", round(d["e2e"]["value"],1))
PY
