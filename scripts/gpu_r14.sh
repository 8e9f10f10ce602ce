bash scripts/gpu_r12.sh 2>&1 | grep -v "cvt_rows\|fold_\|gemm_generic" | grep "passed\|failed\|rror\|prior\|head\|video_pre\|^[0-9]\|assert"
echo "== 1 CTA/SM video_pre"
XATTN_VPRE_1CTA=1 bash scripts/gpu_ll.sh 2>&1 | grep "video_pre"
