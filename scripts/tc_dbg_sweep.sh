for d in 0 1 4 8 5 12 13; do echo "== dbg $d"; SODA_TC_DBG=$d timeout 100 python scripts/conv_probe.py --nocheck "$@" 2>&1 | grep -v Warn; done
