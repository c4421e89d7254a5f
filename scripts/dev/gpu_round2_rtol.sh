#!/bin/bash
# Improvement-based stop of the cooperative Newton solve (JACO_SOLVER_RTOL): timing on the aged roll-out and the GPU parity suite per variant
O=gpurun_out
cp mujoco_jaco_b200/libjaco_b200.so /tmp/prod.so
for r in both; do
  L=mujoco_jaco_b200/libjaco_b200_rtol_$r.so
  echo "== rtol $r"
  python scripts/place_ab.py $L 65536 200 20 -1 2>&1 | cut -c1-60,150-400 | tee -a $O/rtol_ab.log
  python scripts/block_times.py $L 65536 200 -1 2>&1 | tee -a $O/rtol_ab.log
  cp $L mujoco_jaco_b200/libjaco_b200.so
  python -m pytest tests -m gpu -x -q -s > $O/rtol_gputest_$r.log 2>&1; echo "pytest rc=$?" | tee -a $O/rtol_ab.log; tail -2 $O/rtol_gputest_$r.log
  grep -n "cap counters\|excluded from the exact" $O/rtol_gputest_$r.log | head -4
done
cp /tmp/prod.so mujoco_jaco_b200/libjaco_b200.so
