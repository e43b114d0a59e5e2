#!/bin/bash
# The one GPU-box runner (replaces the per-experiment scripts of round 1):  tools/gpu_run.sh <tag> <stage> [<stage> ...]
# Every stage writes gpurun_out/<tag>_<stage>.{log,json,csv}; nothing here is timed under a profiler except the ncu stages.
# Stages:  tests | tests_quick | smoke | bench | bench_short | bench_s7_4 | bench_soup10m | closest_hit | launches | ncu_trace |
#          ncu_shade | ncu_soup10m | ncu_s7_4 | ncu_frame | variants (needs VARIANTS="a.so b.so", SPP) | multi (needs NGPU) | scale (needs NGPU) |
#          build_time (host / GPU builder at 1 M and 10 M) | coherence (ray ordering experiment) | ab_fuse (fused texel records on / off)
set -u
TAG=$1; shift
OUT=gpurun_out; mkdir -p $OUT
PY=python
NGPU=${NGPU:-1}
SPP=${SPP:-64}
trun() { $PY -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "${@:2}"; }
for stage in "$@"; do
  echo "=== $TAG $stage $(date +%T)"
  case $stage in
    tests)        timeout 1500 $PY -m pytest tests -m gpu -x -q --durations=15 > $OUT/${TAG}_tests.log 2>&1; tail -4 $OUT/${TAG}_tests.log ;;
    tests_quick)  timeout 900 $PY -m pytest tests -m gpu -x -q --durations=10 -k "${TESTS_K:-dist or baseline}" > $OUT/${TAG}_tests_quick.log 2>&1; tail -4 $OUT/${TAG}_tests_quick.log ;;
    smoke)        timeout 300 $PY -c 'import __graft_entry__ as g; g.smoke()' > $OUT/${TAG}_smoke.log 2>&1; tail -2 $OUT/${TAG}_smoke.log ;;
    bench)        timeout 900 $PY bench.py --steps ${STEPS:-3} --warmup 3 > $OUT/${TAG}_bench_full.json 2> $OUT/${TAG}_bench_full.err; cut -c1-400 $OUT/${TAG}_bench_full.json ;;
    bench_ref)    timeout 900 $PY bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; cut -c1-300 $OUT/${TAG}_bench_reference.json ;;
    bench_short)  timeout 600 $PY bench.py --steps 3 --warmup 3 --spp $SPP --no-cpu-baseline --no-e2e > $OUT/${TAG}_bench_spp$SPP.json 2> $OUT/${TAG}_bench_spp$SPP.err; cut -c1-300 $OUT/${TAG}_bench_spp$SPP.json ;;
    bench_s7_4)   timeout 600 $PY bench.py --workload s7_4 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_s7_4.json 2> $OUT/${TAG}_bench_s7_4.err; cut -c1-300 $OUT/${TAG}_bench_s7_4.json ;;
    bench_soup10m) timeout 900 $PY bench.py --workload soup_10m --steps 3 --warmup 3 --cpu-sample 480x270x1 > $OUT/${TAG}_bench_soup10m.json 2> $OUT/${TAG}_bench_soup10m.err; cut -c1-400 $OUT/${TAG}_bench_soup10m.json ;;
    closest_hit)  if [ "$NGPU" -gt 1 ]; then trun $NGPU tools/bench_closest_hit.py --rays 64000000 > $OUT/${TAG}_closest_hit_n$NGPU.jsonl 2> $OUT/${TAG}_closest_hit_n$NGPU.err; cat $OUT/${TAG}_closest_hit_n$NGPU.jsonl
                  else timeout 600 $PY tools/bench_closest_hit.py --rays 64000000 > $OUT/${TAG}_closest_hit.jsonl 2> $OUT/${TAG}_closest_hit.err; cat $OUT/${TAG}_closest_hit.jsonl; fi ;;
    launches)     timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches_bench.csv \
                     $PY bench.py --steps 1 --warmup 3 --spp 16 --no-cpu-baseline --no-e2e > $OUT/${TAG}_launches.log 2>&1; $PY tools/launch_share.py $OUT/${TAG}_launches_bench.csv > $OUT/${TAG}_launch_share.txt; cat $OUT/${TAG}_launch_share.txt ;;
    ncu_trace)    timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 200 -c 1 -o $OUT/${TAG}_k_wf_trace -f \
                     $PY bench.py --steps 1 --warmup 3 --spp 32 --no-cpu-baseline --no-e2e > $OUT/${TAG}_ncu_trace.log 2>&1; ls -la $OUT/${TAG}_k_wf_trace.ncu-rep ;;
    ncu_shade)    timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 200 -c 1 -o $OUT/${TAG}_k_wf_shade -f \
                     $PY bench.py --steps 1 --warmup 3 --spp 32 --no-cpu-baseline --no-e2e > $OUT/${TAG}_ncu_shade.log 2>&1; ls -la $OUT/${TAG}_k_wf_shade.ncu-rep ;;
    ncu_soup10m)  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 40 -c 1 -o $OUT/${TAG}_soup10m_k_wf_trace -f \
                     $PY bench.py --workload soup_10m --steps 1 --warmup 3 --spp 4 --no-cpu-baseline --no-e2e > $OUT/${TAG}_ncu_soup10m.log 2>&1; ls -la $OUT/${TAG}_soup10m_k_wf_trace.ncu-rep ;;
    ncu_s7_4)     timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 200 -c 1 -o $OUT/${TAG}_s7_4_k_wf_trace -f \
                     $PY bench.py --workload s7_4 --steps 1 --warmup 3 --spp 32 --no-cpu-baseline --no-e2e > $OUT/${TAG}_ncu_s7_4.log 2>&1; ls -la $OUT/${TAG}_s7_4_k_wf_trace.ncu-rep ;;
    ncu_frame)    # steady-state launch of the roofline leg's shape (one slot group): WORKLOAD, NCU_SPP, NCU_SKIP, KERNEL
                  W=${WORKLOAD:-sponza}; K=${KERNEL:-k_wf_trace}
                  timeout 900 $PY tools/ncu_frame.py --workload $W --spp ${NCU_SPP:-8} > $OUT/${TAG}_frame_${W}.json 2>/dev/null; cat $OUT/${TAG}_frame_${W}.json
                  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s ${NCU_SKIP:-20} -c 1 -o $OUT/${TAG}_${W}_$K -f \
                     $PY tools/ncu_frame.py --workload $W --spp ${NCU_SPP:-8} > $OUT/${TAG}_ncu_frame_${W}_$K.log 2>&1; ls -la $OUT/${TAG}_${W}_$K.ncu-rep ;;
    variants)     # A/B of library variants built with `make OBJ=.. LIB=..`: RTB_LIB=<file> selects the library the binding loads
                  for v in $VARIANTS; do
                    echo "--- $v"
                    RTB_LIB=$v timeout 600 $PY bench.py --steps 3 --warmup 3 --spp $SPP --no-cpu-baseline --no-e2e 2>/dev/null | $PY -c 'import sys,json; d=json.loads(sys.stdin.read()); r=d["roofline"]; print(json.dumps({"value": d["value"], "trace_ms": r["kernel_ms_total"], "other_ms": r["other_kernels_ms_total"], "kernel_mrays_per_s": r["kernel_mrays_per_s"]}))'
                  done > $OUT/${TAG}_variants_spp$SPP.txt 2>&1; cat $OUT/${TAG}_variants_spp$SPP.txt ;;
    multi)        timeout 600 $PY -m pytest tests/test_gpu_dist.py -m gpu -x -q > $OUT/${TAG}_tests_dist_n$NGPU.log 2>&1; tail -3 $OUT/${TAG}_tests_dist_n$NGPU.log
                  timeout 600 $PY -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) tools/check_multigpu.py --full > $OUT/${TAG}_check_multigpu_n$NGPU.log 2>&1; tail -2 $OUT/${TAG}_check_multigpu_n$NGPU.log ;;
    target)       # north_star's target sentence through the CLI (a thread per GPU), image checked against the one-GPU frame
                  timeout 900 $PY tools/target_run.py --spp ${TARGET_SPP:-1024} --out $OUT/${TAG}_target_run_n$NGPU.json > /dev/null 2> $OUT/${TAG}_target_run_n$NGPU.err; cut -c1-600 $OUT/${TAG}_target_run_n$NGPU.json ;;
    bench_auto)   # the driver's command at N GPUs: tile split headline + sample split secondary
                  timeout 900 $PY -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $NGPU --steps ${STEPS:-3} --warmup 3 > $OUT/${TAG}_bench_n${NGPU}_auto.json 2> $OUT/${TAG}_bench_n${NGPU}_auto.err; cut -c1-300 $OUT/${TAG}_bench_n${NGPU}_auto.json ;;
    scale)        for split in sample tile; do
                    trun $NGPU bench.py --gpus $NGPU --steps ${STEPS:-3} --warmup 3 --split $split > $OUT/${TAG}_bench_n${NGPU}_$split.json 2> $OUT/${TAG}_bench_n${NGPU}_$split.err; cut -c1-300 $OUT/${TAG}_bench_n${NGPU}_$split.json
                  done ;;
    scale_soup10m) SP=${SPLIT:-sample}; trun $NGPU bench.py --gpus $NGPU --workload soup_10m --steps ${STEPS:-3} --warmup 3 --split $SP > $OUT/${TAG}_bench_soup10m_n${NGPU}_$SP.json 2> $OUT/${TAG}_bench_soup10m_n${NGPU}_$SP.err; cut -c1-300 $OUT/${TAG}_bench_soup10m_n${NGPU}_$SP.json ;;
    build_time)   timeout 900 $PY tools/build_time.py > $OUT/${TAG}_build_time.jsonl 2> $OUT/${TAG}_build_time.err; cut -c1-200 $OUT/${TAG}_build_time.jsonl ;;
    coherence)    timeout 600 $PY tools/exp_ray_coherence.py > $OUT/${TAG}_ray_coherence.jsonl 2> $OUT/${TAG}_ray_coherence.err; cut -c1-200 $OUT/${TAG}_ray_coherence.jsonl ;;
    ab_fuse)      tools/ab_fuse.sh > $OUT/${TAG}_fuse_ab.txt 2>&1; cat $OUT/${TAG}_fuse_ab.txt ;;
    *) echo "unknown stage $stage" ;;
  esac
done
echo "=== done $(date +%T)"
