# per-role dependency wait + early B: full GPU suite, timelines, per-projection timings, bench
cd /root/repo
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
{
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 1920 --case wo,qkvg
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 640 --case wo_split,w13
python tools/gemm_bench.py --m 1920 640
} > gpurun_out/r2_trace17.log 2>&1
grep -v "stream-K head" gpurun_out/r2_trace17.log
python bench.py --no-config3 > gpurun_out/r2_bench17.json 2> gpurun_out/r2_bench17.err; cut -c1-260 gpurun_out/r2_bench17.json
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench17.json') if l.startswith('{')][0])
print(d['value'], d['ms_per_step'], d['clocks'], d['step_latency_ms'], d['breakdown_ms'], d['generation_roofline']['frac'])
PY
