set -x
python -m pytest tests/test_gpu_bf16.py -q -s -x -k "unsupported or full_batch" 2>&1 | grep -v "^$" | tail -30
python scripts/prof_fwd.py 4096 3 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"v2a_attn|a2v_attn|gemm_tc|video_update" -s 5 -c 5 -o gpurun_out/r3_prof python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log; ls -la gpurun_out/
