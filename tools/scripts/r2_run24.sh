# fused norm tail (then_norm): kernel tests, full suite, A/B of the layer sequence and the bench with ECHO_FUSED_NORM=0/1
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -k "then_norm or split_k or plain" 2>&1 | tail -5
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for f in 0 1; do
  ECHO_FUSED_NORM=$f python bench.py --no-config3 > gpurun_out/r2_bench24_$f.json 2> gpurun_out/r2_bench24_$f.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_bench24_$f.json') if l.startswith('{')][0])
print("ECHO_FUSED_NORM=$f", round(d['value'],2), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], d['step_latency_ms']['cfg_step'], d['step_latency_ms']['plain_step'], round(d['generation_roofline']['frac'],4), d['gpu_launches'])
print(d['config5_streaming']['incremental_prefix']['total_ms'], d['config5_streaming']['incremental_prefix']['first_audio_ms'])
PY
done
