for d in 0 1 2 4 6 8 16 7 31; do echo "dbg=$d"; NPFK_ATTN_DBG=$d timeout 100 python tools/mha_perf.py fast 2>&1 | grep -E "^bench"; done
