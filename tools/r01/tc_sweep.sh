for dbg in 0 16 32 64; do echo "== dbg $dbg"; THA_TC_DBG=$dbg timeout 100 python tools/test_gemm_tc.py 2>&1 | grep "us " | sed -n '2p;4p;6p;7p'; done
