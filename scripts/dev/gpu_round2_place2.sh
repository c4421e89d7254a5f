L=mujoco_jaco_b200/libjaco_b200_diag.so
python scripts/place_ab.py $L 65536 200 20 -1 > gpurun_out/place_ab4_div0.log 2>&1; cat gpurun_out/place_ab4_div0.log
for d in 8 16 32 64; do JACO_KEY_ITER_DIV=$d python scripts/place_ab.py $L 65536 200 20 -1 > gpurun_out/place_ab4_div$d.log 2>&1; echo "div $d"; cat gpurun_out/place_ab4_div$d.log; done
JACO_KEY_ITER_DIV=16 JACO_SORT_TARGET=96 python scripts/place_ab.py $L 65536 200 20 -1 > gpurun_out/place_ab4_div16_t96.log 2>&1; echo "div 16 target 96"; cat gpurun_out/place_ab4_div16_t96.log
JACO_SORT_TARGET=24 python scripts/place_ab.py $L 65536 200 20 -1 > gpurun_out/place_ab4_t24.log 2>&1; echo "div 0 target 24"; cat gpurun_out/place_ab4_t24.log
