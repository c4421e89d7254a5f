#!/bin/bash
# Final round-2 evidence: GPU parity suite, bench lines (default protocol, reference arm, rounds 1-2 action bank, config 5), ncu launch
# list (--cache-control none) and one ncu --set full capture of the step kernel, render bench.  Outputs under gpurun_out/.
O=gpurun_out
python -m pytest tests -m gpu -x -q -s > $O/r02_final_gputest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r02_final_gputest.log
python bench.py > $O/r02_final_bench.json 2> $O/r02_final_bench.err; echo "bench rc=$?"
python bench.py --impl reference > $O/r02_final_bench_reference.json 2> $O/r02_final_bench_reference.err; echo "reference rc=$?"; head -c 400 $O/r02_final_bench_reference.json; echo
python bench.py --action-bank 32 --no-cpu-baseline --no-contact-rich > $O/r02_final_bench_bank32.json 2> $O/r02_final_bench_bank32.err; echo "bank32 rc=$?"
python bench.py --config 5 --no-cpu-baseline > $O/r02_final_bench_config5.json 2> $O/r02_final_bench_config5.err; echo "config5 rc=$?"
python scripts/render_bench.py 256 10 > $O/r02_render_bench.json 2> $O/render_bench.err; echo "render rc=$?"
python - <<'P'
import json
for f in ("r02_final_bench", "r02_final_bench_bank32", "r02_final_bench_config5"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, "value", round(d["value"]), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"]), "launches", d["gpu_launches"], "clocks", d.get("clocks"),
              "roofline", round(d["roofline"]["frac"], 4), "cpu", (d.get("cpu_baseline") or {}).get("value"), "contact_rich", (d.get("contact_rich") or {}).get("value"))
        print("   ", {k: d["episode_stats"][k] for k in ("contact_fraction", "stageB_fraction", "pinv_fraction", "termination_causes")})
    except Exception as e:
        print(f, "unreadable", e)
print(open("gpurun_out/r02_render_bench.json").read()[:300])
P
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:jaco_ --launch-skip 612 -c 18 --csv --log-file $O/r02_final_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-contact-rich > $O/ncu_launches.log 2>&1; echo "ncu list rc=$?"
python - <<'P'
import csv
rows = [r for r in csv.reader(l for l in open('gpurun_out/r02_final_launches.csv') if l.startswith('"'))]
h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value')
for r in rows[1:]: print(r[ki][:40], float(r[vi].replace(',', '')) / 1e6, 'ms')
P
timeout 1500 ncu --set full --metrics smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__issue_active.avg.pct,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum --clock-control none --import-source on -k regex:jaco_step_coop_kernel --launch-skip 206 -c 1 -f -o $O/r02_final_step python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-contact-rich > $O/ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/ncu_summary.py $O/r02_final_step.ncu-rep 65536 $O/r02_final_ncu_step_kernel.json > $O/ncu_summary.log 2>&1; echo "summary rc=$?"; head -c 1500 $O/r02_final_ncu_step_kernel.json
ls -la $O/r02_final_step.ncu-rep
python scripts/ncu_by_function.py $O/r02_final_step.ncu-rep > $O/r02_final_ncu_step_kernel_by_function.txt 2> $O/ncu_by_function.err; head -30 $O/r02_final_ncu_step_kernel_by_function.txt
python scripts/block_times.py mujoco_jaco_b200/libjaco_b200_diag.so 65536 200 -1 0 > $O/r02_final_block_times.log 2>&1; cat $O/r02_final_block_times.log
[ "$(stat -c %s $O/r02_final_step.ncu-rep 2>/dev/null || echo 0)" -gt 40000000 ] && rm -f $O/r02_final_step.ncu-rep
echo done
