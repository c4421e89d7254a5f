"""Aggregate an ncu --set full --import-source on report by source function (inlined code is attributed to the
function whose body holds the source line).  usage: python scripts/ncu_by_function.py report.ncu-rep [kernel-regex]"""
import csv, io, re, subprocess, sys, collections, os

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
fn_re = re.compile(r"^(?:template\s*<[^>]*>\s*)?(?:JHD|static|__global__|__device__|JHD JNOINLINE|inline|extern)[^;(]*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\(")
def functions(path):
    res = []
    try:
        for n, l in enumerate(open(path), 1):
            m = fn_re.match(l)
            if m and not l.rstrip().endswith(";"):
                res.append((n, m.group(1)))
    except OSError:
        pass
    return res
agg = collections.defaultdict(lambda: [0, 0, 0, 0])   # inst, thread inst, samples, no_inst samples
cur, fns, hdr = None, [], None
for row in csv.reader(io.StringIO(out)):
    if not row: continue
    if row[0] == "File Path": cur = row[1]; fns = functions(cur); continue
    if row[0] == "Line No": hdr = row; continue
    if hdr is None or cur is None or not row[0].isdigit(): continue
    ln = int(row[0]); name = "?"
    for n, f in fns:
        if n <= ln: name = f
        else: break
    g = lambda k: float(row[hdr.index(k)] or 0)
    a = agg[(os.path.basename(cur), name)]
    a[0] += g("Instructions Executed"); a[1] += g("Thread Instructions Executed"); a[2] += g("# Samples")
    a[3] += g("stall_no_inst") if "stall_no_inst" in hdr else 0
tot = sum(a[0] for a in agg.values()); tots = sum(a[2] for a in agg.values())
print(f"total warp-instructions {tot:.0f}, samples {tots:.0f}")
print(f"{'file':18s} {'function':28s} {'inst %':>7s} {'thr/inst':>8s} {'samples %':>9s} {'no_inst % of fn':>15s}")
for (f, n), a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:40]:
    print(f"{f:18s} {n:28s} {100*a[0]/tot:7.2f} {a[1]/max(a[0],1):8.1f} {100*a[2]/tots:9.2f} {100*a[3]/max(a[2],1):15.1f}")
