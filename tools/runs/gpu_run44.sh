# ncu --set full of the trace kernel with the 8-entry shared-memory ring (the default build here), same launch as r01m_k_wf_trace
mkdir -p gpurun_out
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/r44_plain.json 2> gpurun_out/r44_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 14 -c 1 -o gpurun_out/r01m_k_wf_trace_ring8 $B > gpurun_out/ncu_trace.log 2>&1; tail -1 gpurun_out/ncu_trace.log
