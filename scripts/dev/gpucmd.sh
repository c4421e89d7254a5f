python scripts/time_variants.py --age 0 main; python scripts/time_variants.py --age 200 main
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 620 -c 12 --csv --log-file gpurun_out/launches_tp.csv python bench.py --steps 3 --warmup 3 --age-steps 200 --no-cpu-baseline > /dev/null 2>&1
python - <<'P'
import csv
rows=[r for r in csv.reader(l for l in open('gpurun_out/launches_tp.csv') if l.startswith('"'))]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value')
for r in rows[1:]: print(r[ki][:40], float(r[vi].replace(',',''))/1e6, 'ms')
P
python scripts/dirty_stats.py 2>&1 | grep "age   0\|age  20\|age 100\|age 200"
