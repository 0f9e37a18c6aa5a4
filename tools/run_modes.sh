for m in 0 1 4 16 20 2 6; do echo "mode $m"; SEVENN_TC_MMA_MODE=$m timeout 100 python tools/bench_gemm.py tc 2>&1 | sed -n 2p; done
