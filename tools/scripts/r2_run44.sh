cd /root/repo
timeout 900 python -m pytest tests/test_gpu_attention_tail.py -x -q 2>&1 | tail -2
for lib in old new old new; do
  echo "== attention kernel: $lib"
  if [ $lib = old ]; then export ECHO_B200_LIB=old; else unset ECHO_B200_LIB; fi
  python tools/gemm_bench.py --m 1920 640 --only attn,seq | grep -v layer
done
unset ECHO_B200_LIB
ECHO_B200_LIB=debug python tools/attn_trace.py --rows 3 2>&1 | head -7
ECHO_B200_LIB=debug python tools/attn_trace.py --rows 1 2>&1 | head -7
