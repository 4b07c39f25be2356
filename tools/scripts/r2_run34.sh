cd /root/repo
timeout 900 python -m pytest tests/test_gpu_attention_tail.py tests/test_gpu_streaming.py -x -q 2>&1 | tail -2
ECHO_B200_LIB=debug ECHO_ATTN_SPLITS=4 python tools/attn_trace.py --rows 1 --seq 160 --prefix 480 2>&1 | sed -n '1p;6p'
ECHO_B200_LIB=debug python tools/attn_trace.py --rows 3 --seq 160 --prefix 480 2>&1 | sed -n '1p;6p'
ECHO_B200_LIB=debug ECHO_ATTN_SPLITS=1 python tools/attn_trace.py --rows 3 --seq 160 --prefix 480 2>&1 | sed -n '1p'
for sp in 1 0; do echo "== ECHO_ATTN_SPLITS=$sp (0 = automatic)"; ECHO_ATTN_SPLITS=$sp python tools/gemm_bench.py --m 160 480 --seq 160 --only attn,seq | grep -v layer; done
for sp in 1 0; do
ECHO_ATTN_SPLITS=$sp python bench.py --no-config3 --no-cfg-split > gpurun_out/r2_bench34_$sp.json 2> gpurun_out/r2_bench34_$sp.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_bench34_$sp.json') if l.startswith('{')][0])
print("ECHO_ATTN_SPLITS=$sp", round(d['value'],2), d['step_latency_ms']['plain_step'], d['config5_streaming']['incremental_prefix']['total_ms'], d['config5_streaming']['incremental_prefix']['first_audio_ms'])
PY
done
