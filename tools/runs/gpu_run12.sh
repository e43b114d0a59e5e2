mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --width 960 --height 272 --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/bench_$name.json")); r=d["roofline"]; print(round(d["value"],1), round(d["ms_per_step"],1), round(r["kernel_ms_total"],1), round(r["other_kernels_ms_total"],1), round(d["e2e"]["value"],1))
PY
}
run small_rpl1 RTB_TRACE_RPL=1
run small_rpl2 RTB_TRACE_RPL=2
run small_rpl4 RTB_TRACE_RPL=4
run small_rpl8 RTB_TRACE_RPL=8
run small_rpl4_g1 RTB_TRACE_RPL=4 RTB_WF_GROUPS=1
run small_mk RTB_TRACE_RPL=1 
timeout 300 python bench.py --width 960 --height 272 --spp 64 --steps 3 --warmup 2 --no-cpu-baseline --megakernel 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('megakernel small', round(d['value'],1), round(d['ms_per_step'],1))"
