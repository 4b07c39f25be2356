# multi-GPU evidence at N = 8 (data-parallel prompts, configs[2], CFG-branch split on ranks 0-2)
cd /root/repo
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err
tail -2 gpurun_out/r2_bench_n8.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_n8.json') if l.startswith('{')][0])
print(d['value'], d['ms_per_step'], d['n_gpus'], d['clocks'])
print(json.dumps(d.get('config3'))[:300])
print(json.dumps(d.get('cfg_split'))[:300])
PY
