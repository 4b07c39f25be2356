cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for e in 0 1; do
ECHO_ENCODER_STREAMS=$e python bench.py --no-config3 --no-cfg-split > gpurun_out/r2_bench39_$e.json 2> gpurun_out/r2_bench39_$e.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_bench39_$e.json') if l.startswith('{')][0])
print("ECHO_ENCODER_STREAMS=$e", round(d['value'],2), round(d['ms_per_step'],3), d['step_latency_ms']['cfg_step'], d['step_latency_ms']['plain_step'], d['breakdown_ms'], d['config5_streaming']['incremental_prefix']['total_ms'], d['config5_streaming']['incremental_prefix']['first_audio_ms'])
PY
done
