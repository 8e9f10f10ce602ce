set -x
python scripts/prof_fwd.py 4096 3 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"v2a_attn|a2v_attn|head_pooled|video_update|video_pre" -s 5 -c 5 -o gpurun_out/r5_prof python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
