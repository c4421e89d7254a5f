L=mujoco_jaco_b200/libjaco_b200_diag.so
python scripts/block_times.py $L 65536 200 -1 > gpurun_out/block_times2.log 2>&1; cat gpurun_out/block_times2.log
for ta in 6 12 24; do for tb in 48 96; do echo "targetA $ta targetB $tb"; JACO_SORT_TARGET_A=$ta JACO_SORT_TARGET=$tb python scripts/place_ab.py $L 65536 200 20 -1 2>&1 | cut -c1-60,150-330 | tee -a gpurun_out/place_ab5.log; done; done
echo "hi 256"; JACO_SORT_HI=256 python scripts/place_ab.py $L 65536 200 20 -1 2>&1 | cut -c1-60,150-330 | tee -a gpurun_out/place_ab5.log
JACO_SORT_TARGET_A=6 python scripts/block_times.py $L 65536 200 -1 > gpurun_out/block_times3.log 2>&1; cat gpurun_out/block_times3.log
