for d in 0 1; do
echo "== dbg $d"
XATTN_LNRES_DBG=$d timeout -k 5 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gemm_tc_kernel --csv --log-file gpurun_out/ll_$d.csv python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
python scripts/launch_summary.py gpurun_out/ll_$d.csv | grep "4, 6"
done
