cd /root/repo
timeout 900 python -m pytest tests/test_gpu_attention_tail.py -x -q 2>&1 | tail -2
export ECHO_B200_LIB=debug
for sp in 1 2 4; do echo "== splits $sp seq 160 prefix 480 rows 1"; ECHO_ATTN_SPLITS=$sp python tools/attn_trace.py --rows 1 --seq 160 --prefix 480 2>&1 | head -1; done
for sp in 1 2 3; do echo "== splits $sp seq 640 rows 1"; ECHO_ATTN_SPLITS=$sp python tools/attn_trace.py --rows 1 2>&1 | head -1; done
echo "== rows 3 seq 160 prefix 480 auto"; python tools/attn_trace.py --rows 3 --seq 160 --prefix 480 2>&1 | head -1
unset ECHO_B200_LIB
for sp in 1 2 3 4; do echo "== ECHO_ATTN_SPLITS=$sp"; ECHO_ATTN_SPLITS=$sp python tools/gemm_bench.py --m 640 --only attn,seq; done
for sp in 1 4; do echo "== ECHO_ATTN_SPLITS=$sp (seq 160)"; ECHO_ATTN_SPLITS=$sp python tools/gemm_bench.py --m 160 480 --seq 160 --only attn,seq; done
