XATTN_TC_TRACE=1 timeout -k 5 120 python scripts/prof_fwd.py 4096 2 2>&1 | tail -40
