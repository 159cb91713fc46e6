# tcgen05 GEMM with stages compiled out (experiment library, make ABLATE=1): TB200_GEMM_DBG bits
# 1 no epilogue body, 2 no 2-bit -> u8 expansion, 4 no MMAs, 8 no activation TMA traffic
for d in ${DBGS:-0 1 2 4 8 3 7 15}; do echo -n "DBG=$d  "; TB200_GEMM_DBG=$d TB200_LIB=ternary_vllm_b200/libternary_b200_ablate.so python tools/gemm_probe.py "$@" 2>&1 | grep path; done
