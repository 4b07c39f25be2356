# evidence: the full GPU suite (log kept under profiles/), the N=1 bench line, both arms
cd /root/repo
python -m pytest tests -m gpu -x -q -s -p no:cacheprovider 2>&1 | grep -v "^$" | tail -150 > gpurun_out/r2_gpu_tests.log
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err
tail -5 gpurun_out/r2_gpu_tests.log; cut -c1-600 gpurun_out/r2_bench_n1.json; cut -c1-400 gpurun_out/r2_bench_ref.json
