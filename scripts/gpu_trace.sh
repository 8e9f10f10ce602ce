XATTN_TC_TRACE=1 timeout -k 5 120 python scripts/prof_fwd.py 4096 1 2>&1 | grep -A 18 "${1:-a2v}_tc trace" | head -20
