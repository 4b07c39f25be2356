cd /root/repo
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py -x -q 2>&1 | tail -2
python - <<'PY'
import torch, time
from echo_tts_b200 import schema as S
from echo_tts_b200.inference import load_model_from_state_dict
model = load_model_from_state_dict(S.synth_state_dict(S.dit_schema(S.ECHO_BASE, blockwise=False), init="torch"))
t = torch.linspace(0.999, 0.0, 41)[:-1].cuda()
for _ in range(3): model.build_mod_table(t)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): model.build_mod_table(t)
e1.record(); torch.cuda.synchronize()
print("build_mod_table (40 timesteps): %.3f ms" % (e0.elapsed_time(e1) / 10))
PY
python bench.py --no-config3 --no-cfg-split > gpurun_out/r2_bench37.json 2> gpurun_out/r2_bench37.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench37.json') if l.startswith('{')][0])
print(round(d['value'],2), round(d['ms_per_step'],2), d['clocks']['sm_mhz'], d['step_latency_ms'], d['breakdown_ms'], round(d['generation_roofline']['frac'],4))
print(d['config5_streaming']['incremental_prefix']['total_ms'], d['config5_streaming']['incremental_prefix']['first_audio_ms'])
PY
