cd /root/repo
timeout 900 python -m pytest tests/test_gpu_attention_tail.py -x -q 2>&1 | tail -3
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/config_bench.py 2>&1 | tail -8
