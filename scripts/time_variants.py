"""Time bench.py's device step for several builds of the library (scripts/build_variant.sh).
usage: python scripts/time_variants.py tag1 tag2 ...   ("main" = the shipped libjaco_b200.so)"""
import json, os, subprocess, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for tag in sys.argv[1:]:
    lib = os.path.join(REPO, "mujoco_jaco_b200", "libjaco_b200.so" if tag == "main" else f"libjaco_b200_{tag}.so")
    if not os.path.exists(lib):
        print(tag, "missing", flush=True); continue
    env = dict(os.environ, JACO_B200_LIB=lib)
    p = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--steps", "8", "--warmup", "6", "--no-cpu-baseline"],
                       env=env, capture_output=True, text=True)
    try:
        d = json.loads(p.stdout.strip().splitlines()[-1])
        print(tag, "ms_per_step %.2f" % d["ms_per_step"], "value %.0f" % d["value"], "e2e %.0f" % d["e2e"]["value"], flush=True)
    except Exception as e:
        print(tag, "failed", p.stderr[-500:], flush=True)
