# one more default bench line of the round's final kernels on a fresh box (box-to-box range of the headline)
cd /root/repo
python bench.py > gpurun_out/r2_bench_n1_b.json 2> gpurun_out/r2_bench_n1_b.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_n1_b.json') if l.startswith('{')][0])
print(d['value'], d['ms_per_step'], d['clocks'], d['breakdown_ms'], d['step_latency_ms']['cfg_step'], d['step_latency_ms']['plain_step'], d['roofline']['frac'], d['generation_roofline']['frac'])
PY
