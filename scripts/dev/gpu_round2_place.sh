python scripts/place_ab.py mujoco_jaco_b200/libjaco_b200_diag.so 65536 200 20 0 64 -1 > gpurun_out/place_ab3.log 2>&1; echo "ab rc=$?"; cat gpurun_out/place_ab3.log
JACO_SORT_TARGET=32 python scripts/place_ab.py mujoco_jaco_b200/libjaco_b200_diag.so 65536 200 20 -1 > gpurun_out/place_ab3b.log 2>&1; cat gpurun_out/place_ab3b.log
JACO_SORT_TARGET=80 python scripts/place_ab.py mujoco_jaco_b200/libjaco_b200_diag.so 65536 200 20 -1 > gpurun_out/place_ab3c.log 2>&1; cat gpurun_out/place_ab3c.log
python -m pytest tests -m gpu -x -q > gpurun_out/gputest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/gputest.log
python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-contact-rich > gpurun_out/bench_iid.json 2> gpurun_out/bench_iid.err; echo "bench rc=$?"; head -c 1500 gpurun_out/bench_iid.json
python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-contact-rich --action-bank 32 > gpurun_out/bench_bank32.json 2> gpurun_out/bench_bank32.err; echo "bench rc=$?"; head -c 400 gpurun_out/bench_bank32.json
