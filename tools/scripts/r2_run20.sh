# split-KV attention, fast division in the GEMM scheduler, 8-way split-K for one-row-tile forwards
cd /root/repo
python -m pytest tests/test_gpu_attention_tail.py tests/test_gpu_kernels.py -x -q 2>&1 | tail -3
{
for s in 1 2 3 4; do echo "== ECHO_ATTN_SPLITS=$s"; ECHO_ATTN_SPLITS=$s python tools/gemm_bench.py --m 640 --only attn,seq; done
ECHO_ATTN_SPLITS=2 ECHO_B200_LIB=debug python tools/attn_trace.py --rows 1 | head -8
} > gpurun_out/r2_attn20.log 2>&1
cat gpurun_out/r2_attn20.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --no-config3 > gpurun_out/r2_bench20.json 2> gpurun_out/r2_bench20.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench20.json') if l.startswith('{')][0])
print(d['value'], d['ms_per_step'], d['clocks'], d['step_latency_ms'], d['breakdown_ms'], d['generation_roofline']['frac'])
print(d['config5_streaming']['incremental_prefix'])
PY
