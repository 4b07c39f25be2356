# multi-GPU evidence at N = 4 with the round's final kernels (data-parallel prompts, configs[2], CFG-branch split on ranks 0-2)
cd /root/repo
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/r2_bench_n4.json 2> gpurun_out/r2_bench_n4.err
tail -2 gpurun_out/r2_bench_n4.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_n4.json') if l.startswith('{')][0])
print(d['value'], d['ms_per_step'], d['n_gpus'], d['clocks'])
print(json.dumps(d.get('config3'))[:300])
print(json.dumps(d.get('cfg_split'))[:400])
PY
