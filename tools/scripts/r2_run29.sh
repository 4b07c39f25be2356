cd /root/repo
timeout 900 python -m pytest tests/test_gpu_attention_tail.py -x -q 2>&1 | tail -3
python tools/gemm_bench.py --m 1920 640 --only attn,wo,w13,seq
ECHO_B200_LIB=debug python tools/attn_trace.py --rows 3 2>&1 | head -12
ECHO_B200_LIB=debug python tools/attn_trace.py --rows 1 2>&1 | head -5
timeout 900 python -m pytest tests/test_gpu_full_size.py tests/test_gpu_parity.py tests/test_gpu_streaming.py -x -q 2>&1 | tail -3
