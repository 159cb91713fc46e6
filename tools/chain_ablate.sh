# chain kernel with parts of the dot-product loop compiled out (experiment library, make ABLATE=1):
# TB200_CHAIN_DBG bits: 1 no IMMAs, 2 no field masks, 4 no weight LDS.  Untraced timing (chain_bench) + optional trace.
for d in ${@:-0 1 4 5}; do echo "DBG=$d"; TB200_CHAIN_DBG=$d TB200_LIB=ternary_vllm_b200/libternary_b200_ablate.so timeout 120 python tools/chain_bench.py --fast --tokens 1 8 2>&1 | cut -c1-120; done
