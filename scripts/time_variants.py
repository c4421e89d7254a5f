"""Time bench.py's device step for several builds of the library (scripts/build_variant.sh).  Development tool.
usage: python scripts/time_variants.py [--age N] tag1 tag2 ...   ("main" = the shipped libjaco_b200.so)"""
import json, os, subprocess, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = sys.argv[1:]
age = "200"
if args and args[0] == "--age":
    age = args[1]; args = args[2:]
for tag in args:
    lib = os.path.join(REPO, "mujoco_jaco_b200", "libjaco_b200.so" if tag == "main" else f"libjaco_b200_{tag}.so")
    if not os.path.exists(lib):
        print(tag, "missing", flush=True); continue
    code = (f"import sys; sys.path.insert(0, {REPO!r}); from mujoco_jaco_b200 import _lib; _lib.SO_PATH = {lib!r}; _lib.needs_build = lambda: False; "
            f"sys.argv = ['bench.py', '--steps', '8', '--warmup', '4', '--no-cpu-baseline', '--age-steps', {age!r}]; import runpy; runpy.run_path({os.path.join(REPO, 'bench.py')!r}, run_name='__main__')")
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    try:
        d = json.loads(p.stdout.strip().splitlines()[-1])
        print(tag, "age", age, "ms_per_step %.2f" % d["ms_per_step"], "value %.0f" % d["value"], "e2e %.0f" % d["e2e"]["value"], flush=True)
    except Exception as e:
        print(tag, "failed", p.stdout[-300:], p.stderr[-800:], flush=True)
