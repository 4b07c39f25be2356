# final evidence of the round: the full GPU suite (log kept under profiles/), the default N=1 bench line, the reference arm
cd /root/repo
python -m pytest tests -m gpu -x -q -s -p no:cacheprovider 2>&1 | grep -v "^$" | tail -170 > gpurun_out/r2_gpu_tests.log
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke.log 2>&1
tail -3 gpurun_out/r2_gpu_tests.log; tail -2 gpurun_out/r2_smoke.log; cut -c1-300 gpurun_out/r2_bench_n1.json; cut -c1-300 gpurun_out/r2_bench_ref.json
