set -x
python bench.py > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; tail -c 600 gpurun_out/r02_bench.err; cat gpurun_out/r02_bench.json | cut -c1-3000
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference.json 2>/dev/null; cat gpurun_out/r02_bench_reference.json | cut -c1-1200
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --launch-skip 410 -c 16 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-contact-rich > /dev/null 2>&1
grep jaco gpurun_out/r02_launches.csv | cut -d, -f5,15 | tail -8
timeout 900 ncu --set full --metrics smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum --clock-control none --import-source on -k regex:jaco_step_coop_kernel --launch-skip 206 -c 1 -f -o gpurun_out/r02_full_coop python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-contact-rich > gpurun_out/ncu_full.log 2>&1; tail -2 gpurun_out/ncu_full.log | cut -c1-200
