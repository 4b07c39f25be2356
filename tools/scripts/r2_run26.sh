cd /root/repo
{ ECHO_B200_LIB=debug python tools/gemm_trace.py --m 1920 --case wo_norm; ECHO_B200_LIB=debug python tools/gemm_trace.py --m 640 --case wo_split_norm,w2_split_norm; ECHO_B200_LIB=debug python tools/gemm_trace.py --m 160 --seq 160 --case wo_split_norm; } 2>&1 | grep "norm tail\|^=="
python -m pytest tests/test_gpu_kernels.py -x -q -k "then_norm" 2>&1 | tail -2
for f in 0 1; do
  ECHO_FUSED_NORM=$f python bench.py --no-config3 > gpurun_out/r2_bench26_$f.json 2> gpurun_out/r2_bench26_$f.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_bench26_$f.json') if l.startswith('{')][0])
print("ECHO_FUSED_NORM=$f", round(d['value'],2), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], d['step_latency_ms']['cfg_step'], d['step_latency_ms']['plain_step'], round(d['generation_roofline']['frac'],4), d['gpu_launches'])
print(d['config5_streaming']['incremental_prefix']['total_ms'], d['config5_streaming']['incremental_prefix']['first_audio_ms'])
PY
done
