for d in 0 1 2 8 11; do echo "== RES_DBG=$d"; B200DP_RES_DBG=$d timeout 100 python tools/kernel_bench.py --arch keras128 --out gpurun_out/kb_dbg.json 2>&1 | grep -i "conv2_fwd\|conv2_dgrad"; done
