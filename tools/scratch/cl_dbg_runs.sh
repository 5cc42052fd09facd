for v in clr; do for d in 0 1 128 4 32 36 8 16 24 64 255; do echo "== $v dbg $d"; B200_CL_DBG=$d timeout 120 python tools/cluster_check.py C2 50 $v 2>&1 | grep "^$v" | cut -c1-110; done; done
