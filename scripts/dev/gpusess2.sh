timeout -s KILL 1800 python -m pytest tests -m gpu -q 2>&1 | tail -3
python bench.py --no-cpu-baseline > gpurun_out/r02_bench_b.json 2> gpurun_out/r02_bench_b.err; tail -c 300 gpurun_out/r02_bench_b.err; python - <<'P'
import json
d=json.load(open('gpurun_out/r02_bench_b.json'))
for k in ("value","ms_per_step","gpu_launches","clocks","contact_rich","episode_stats"): print(k, d[k])
print("e2e", d["e2e"]["value"], "roofline", d["roofline"]["frac"], d["roofline"]["flops_source"][:60])
P
python bench.py --config 5 --no-cpu-baseline > gpurun_out/r02_bench_config5.json 2> gpurun_out/r02_bench_c5.err; tail -c 300 gpurun_out/r02_bench_c5.err; python - <<'P'
import json
d=json.load(open('gpurun_out/r02_bench_config5.json'))
for k in ("metric","value","ms_per_step","gpu_launches","episode_stats"): print(k, d[k])
print("e2e", d["e2e"]["value"], d["config"]["workload"][:80])
P
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --launch-skip 820 -c 12 --csv --log-file gpurun_out/r02_launches_b.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-contact-rich > /dev/null 2>&1
python - <<'P'
import csv
rows=[r for r in csv.reader(l for l in open('gpurun_out/r02_launches_b.csv') if l.startswith('"'))]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value')
for r in rows[1:]: print(r[ki][:40], float(r[vi].replace(',',''))/1e6, 'ms')
P
