cd /root/repo
for c in "2 2" "2 3" "3 3" "3 4" "4 4"; do set -- $c; echo "== M=640 splits wo=$1 w2=$2"; python tools/gemm_bench.py --m 640 --splits-wo $1 --splits-w2 $2 --only seq,wo/s$1,w2/s$2,addnorm | grep -v layer; done
for c in "4 4" "8 8" "4 8" "6 8"; do set -- $c; echo "== M=160 splits wo=$1 w2=$2"; python tools/gemm_bench.py --m 160 --seq 160 --splits-wo $1 --splits-w2 $2 --only seq,wo/s$1,w2/s$2 | grep -v layer; done
for c in "4 4" "3 4" "2 4"; do set -- $c; echo "== M=480 splits wo=$1 w2=$2"; python tools/gemm_bench.py --m 480 --seq 160 --splits-wo $1 --splits-w2 $2 --only seq,wo/s$1,w2/s$2 | grep -v layer; done
