#!/usr/bin/env python
"""bench.py -- Jaco picking env-steps/s on N B200s (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--envs E] [--impl reference]

A "step" is one JacoMujocoEnv.step over the whole batch: 50 fused physics substeps + OSC + obs/reward/done for
every env (config C3 of SURVEY.md §8d: picking, reset distribution, actions U(-1,1)^7, auto-reset).
`value` is device-resident throughput (inputs in HBM); `e2e` is the same metric through the host-buffer C-ABI
call (jaco_step_host: pinned actions H2D, obs/reward/done D2H inside the timed region).
`--impl reference` times the CPU restatement of the reference path (oracle/, one env per process on all host
cores) -- the reference's own mujoco-py path is unavailable on this box.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
_OUT = sys.stdout

# algorithmic bytes per env-step of the step kernel (DESIGN.md "roofline"): SoA state rows read + written
# (23 qpos + 21 qvel + 72 aux floats each way) + action 28 B + obs 104 B + reward 4 + wb 4 + done 1 + succ 1
ALG_BYTES_PER_ENVSTEP = 2 * (23 + 21 + 72) * 4 + 28 + 104 + 4 + 4 + 1 + 1
FLOPS_FALLBACK = 1.9e6   # SURVEY 8d a-priori estimate, used only when no instrumented count is committed


def _ncu_profile():
    """The committed ncu --set full summary of the step kernel on the aged bench workload (profiles/r02_ncu_step_kernel.json):
    dram bytes per launch and the INSTRUMENTED FP32 operation count (hardware counters: predicated-on thread-level
    FADD + FMUL + 2 x FFMA of one launch = one env step of 65 536 envs)."""
    for name in ("r02_final_ncu_step_kernel.json", "r02_ncu_step_kernel.json", "r01_ncu_step_kernel.json"):
        p = os.path.join(REPO, "profiles", name)
        try:
            d = json.load(open(p)); d["file"] = "profiles/" + name
            return d
        except Exception:
            continue
    return {}


def _peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.reasons, self.stop_flag, self.max_mhz = gpu_index, [], set(), False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0])); self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if "Active" in v and "Not" not in v:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def _oracle_worker(args):
    wid, task, seconds, seed, warm, fast = args
    import numpy as np
    from oracle.oracle import OracleEnv
    env = OracleEnv(task=task, seed=seed, env_id=wid, fast=fast)
    env.reset()
    rng = np.random.default_rng(wid)
    for _ in range(warm):                      # 200 warm-up env steps per worker (ages the episode like the GPU arm)
        if env.step(rng.uniform(-1, 1, 7))[2]:
            env.reset()
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        _, _, d, _, _ = env.step(rng.uniform(-1, 1, 7))
        n += 1
        if d:
            env.reset()
    return n, time.perf_counter() - t0


def cpu_baseline(seconds, task="picking", workers=None, warm=200):
    """Oracle (CPU restatement of the reference path), one env per process, all host cores, -O3 -march=native build made on
    this host (falls back to the portable -O2 build when that fails)."""
    import multiprocessing as mp
    from oracle import oracle as _o
    _o.build()
    fast = _o.build_fast()
    workers = workers or os.cpu_count() or 1
    with mp.get_context("spawn").Pool(workers) as pool:
        res = pool.map(_oracle_worker, [(w, task, seconds, 0, warm, fast) for w in range(workers)])
    rate = sum(n / t for n, t in res)
    return {"value": rate, "unit": "env-steps/s", "cores": workers, "kind": "port",
            "ms_per_env_step_per_core": 1e3 * workers / rate if rate > 0 else None,
            "build": "gcc -O3 -march=native" if fast else "gcc -O2 (portable; the native build failed)",
            "sample": f"{workers} processes x 1 env, picking roll-out (reset distribution, U(-1,1)^7 actions, auto-reset), {warm} warm-up env steps then "
                      f"{seconds:.0f} s wall each; oracle/jaco_oracle.c float64 (CPU restatement; mujoco-py unavailable on this box)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step_s = max(2.0, min(10.0, 60.0 / max(1, args.steps + args.warmup)))
    vals = []
    for i in range(args.warmup + args.steps):
        cb = cpu_baseline(per_step_s, warm=200 if i == 0 else 20)
        if i >= args.warmup:
            vals.append(cb["value"])
    v = sum(vals) / len(vals)
    cb["value"] = v
    cb["ms_per_env_step_per_core"] = 1e3 * cb["cores"] / v
    print(file=_OUT, flush=True, *[json.dumps({
        "impl": "reference", "metric": "jaco_picking_env_steps_per_sec", "value": v, "unit": "env-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * cb["cores"] / v,   # wall time of one env step of one env on one core (the reference is one env per process)
        "sample_window_s": per_step_s,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "picking roll-out, 1 env per host process (CPU restatement of JacoMujocoEnv.step)"},
        "cpu_baseline": cb, "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--envs", type=int, default=65536, help="envs per GPU (BASELINE config 3: 65536)")
    ap.add_argument("--task", default="picking")
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--age-steps", type=int, default=200,
                    help="untimed roll-out steps before the warm-up (SURVEY 8d C3: 200): the workload drifts as episodes age "
                         "(more arm-scene contacts), so timing straight after reset() would measure an easier mix")
    ap.add_argument("--action-bank", type=int, default=0,
                    help="0 (default): a fresh U(-1,1)^7 action batch is drawn on the device for every step, outside the timed region "
                         "(BASELINE config 3 'actions U(-1,1)^7', what the CPU arm does too); K > 0: cycle through K pre-generated batches "
                         "(rounds 1-2 used K = 32: every env then repeats a 32-step action sequence whose net displacement walks the arm "
                         "into the singular posture, a harder mix than i.i.d. actions)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-contact-rich", action="store_true", help="skip the secondary contact-rich line (SURVEY 8d C3-ii)")
    ap.add_argument("--config", type=int, default=3, choices=[3, 5],
                    help="BASELINE config: 3 = picking roll-out with U(-1,1)^7 actions (headline); 5 = HPC composite policy in the loop "
                         "(16384 envs per GPU, policy forward + env step on the device, random-initialised weights of the saved architecture)")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: libraries that log to fd 1 (NCCL's version banner) are sent to stderr
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from mujoco_jaco_b200 import _lib
    from mujoco_jaco_b200.env import JacoVecEnv, NACT, NOBS
    from mujoco_jaco_b200.shard import gather_per_rank, max_over_ranks, reduce_stats, shard_range

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries exactly one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    if args.config == 5 and args.envs == 65536:
        args.envs = 16384
    n, K, W = args.envs, args.steps, max(3, args.warmup)
    policy = None
    if args.config == 5:
        from mujoco_jaco_b200.policy import CompositePolicy
        policy = CompositePolicy.random_init(seed=0).to(torch.device("cuda", local)).requires_grad_(False)
    env_offset, _ = shard_range(rank, world, n)   # independent shard of the global env range: no data-path collective
    env = JacoVecEnv(n, task=args.task, seed=0, device=local, env_offset=env_offset, auto_reset=True)
    env.reset()
    gen = torch.Generator(device=dev); gen.manual_seed(1234 + rank)
    bank = max(args.action_bank, 0)
    actions = torch.rand(max(bank, 8), n, NACT, device=dev, generator=gen) * 2 - 1   # action bank (and the contact-rich line's actions)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def act(i):
        """the step's action batch on the device: config 3 = pre-generated U(-1,1)^7, config 5 = composite policy forward on the
        env's own observation buffer (no host round trip)"""
        if policy is None:
            if bank:
                return actions[i % bank]
            return torch.rand(n, NACT, device=dev, generator=gen) * 2 - 1     # fresh i.i.d. batch, resident in HBM before the timed region
        with torch.no_grad():
            return policy(env.obs, deterministic=False)[0]

    for i in range(args.age_steps):
        env.step(act(i))
    for i in range(W):
        env.step(act(i))
    env.stats(clear=True)
    sampler = ClockSampler(local)      # every rank samples its own GPU (per_rank table); rank 0's summary is the line's `clocks`
    sampler.start()
    # L2 is flushed between timed iterations (a 256 MB fill, outside the per-step event pairs): the state working set
    # (30 MB) would otherwise stay resident in the 126 MB L2 from one step to the next
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    # `value` (device-resident, CUDA events) and `e2e` (host buffers through jaco_step_host, host clock) are measured on
    # ALTERNATING steps of the same aged population, so that both see the same episode-age mix: step 2i is timed on the device,
    # step 2i+1 end to end.  L2 is flushed before either (outside both timed regions).
    a_hosts = [torch.empty(n, NACT).uniform_(-1, 1).pin_memory() for _ in range(8)]
    outs = (torch.empty(n, NOBS).pin_memory(), torch.empty(n).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory(),
            torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n).pin_memory())
    outs_np = tuple(o.numpy() for o in outs)
    a_nps = [a.numpy() for a in a_hosts]
    env.step_host(a_nps[-1], outs_np)
    env.stats(clear=True)
    launches0 = env.launch_count()
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    e2e_s = 0.0
    barrier()
    for i in range(K):
        flush.fill_(i & 255)
        a_dev = act(W + i) if policy is None else None    # config 3: the action batch exists in HBM before the timed region starts
        ev0[i].record()
        env.step(a_dev if policy is None else act(W + i))   # config 5: the policy forward is inside the timed region
        ev1[i].record()
        flush.fill_((i + 1) & 255)
        torch.cuda.synchronize()
        if policy is None and not bank:
            a_hosts[i % len(a_hosts)].uniform_(-1, 1)        # fresh host actions for this e2e step (pinned, in place), outside the timed region
        t0 = time.perf_counter()
        if policy is None:
            env.step_host(a_nps[i % len(a_nps)], outs_np)   # returns after its own stream synchronisation
        else:   # user-facing loop of main.py:252-260 on host buffers: obs (host) -> predict_subgoal -> action (host) -> step
            a_np, _, _ = policy.predict_subgoal(outs_np[0], deterministic=False)
            env.step_host(a_np, outs_np)
        e2e_s += time.perf_counter() - t0
    barrier()
    kernel_ms = [ev0[i].elapsed_time(ev1[i]) for i in range(K)]
    my_ms = sum(kernel_ms)
    launches = (env.launch_count() - launches0) // 2      # kernels of the K device-timed steps (the K e2e steps launch the same ones)
    my_stats = env.stats(clear=True).clone()
    total_ms = max_over_ranks(my_ms, device=dev)      # device time, max over ranks
    stats = reduce_stats(my_stats.clone())             # the only collective of the path: episode statistics (NCCL)
    value = world * n * K / (total_ms * 1e-3)
    e2e_value = world * n * K / max_over_ranks(e2e_s, device=dev)
    nsub = n * 2 * K * 50                               # env-substeps behind the statistics (device-timed + e2e steps)
    # per-rank table (a gather of a few floats; evidence for the scaling analysis, not a data-path exchange)
    per_rank = gather_per_rank(torch.tensor([my_ms / K, float(my_stats[11]) / nsub, float(my_stats[12]) / nsub,
                                             float(my_stats[13]) / nsub, float(my_stats[0])], device=dev))

    # ---- secondary line, contact-rich mode: the 4096 in-contact snapshots of tests/golden tiled to n envs,
    # re-injected before every timed step (rank 0 only; not the headline)
    contact_rich = None
    snap_path = os.path.join(REPO, "tests", "golden", "contact_snapshots.npz")
    if rank == 0 and not args.no_contact_rich and args.task == "picking" and policy is None and os.path.exists(snap_path):
        Z = np.load(snap_path)
        reps = (n + len(Z["qpos"]) - 1) // len(Z["qpos"])
        tile = lambda x: torch.tensor(np.tile(np.asarray(x, np.float32), (reps, 1))[:n]).to(dev)
        sq, sv, sa = tile(Z["qpos"]), tile(Z["qvel"]), tile(Z["aux"])
        ms = []
        for i in range(6):
            env.set_state(sq, sv, sa)
            env.stats(clear=True)
            flush.fill_(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); env.step(actions[i]); e1.record(); torch.cuda.synchronize()
            if i >= 2:
                ms.append(e0.elapsed_time(e1))
        cs = env.stats(clear=True).cpu().numpy()
        cms = sum(ms) / len(ms)
        contact_rich = {"value": n / (cms * 1e-3), "unit": "env-steps/s", "ms_per_step": cms, "steps": len(ms),
                        "workload": f"4096 in-contact snapshots (tests/golden/contact_snapshots.npz) tiled to {n} envs, re-injected before every step",
                        "contact_fraction": float(cs[13]) / (n * 50), "stageB_fraction": float(cs[11]) / (n * 50),
                        "iteration_cap_hits": float(cs[10]), "contact_cap_hits": float(cs[9])}
    sampler.stop_flag = True
    sampler.join(timeout=2)
    clk = sampler.summary()
    per_rank_hw = gather_per_rank(torch.tensor([min(kernel_ms), max(kernel_ms), float(clk["sm_mhz"] or 0), float(len(clk["reasons"]))], device=dev))

    if rank == 0:
        peaks, which = _peaks()
        L = _lib.load()
        fp32_peak = float(L.jaco_fp32_peak_tflops(local, 5))
        prof = _ncu_profile()
        kms = sum(kernel_ms) / len(kernel_ms)
        flops = float(prof.get("fp32_flops_per_env_step", FLOPS_FALLBACK))
        hbm_ach = ALG_BYTES_PER_ENVSTEP * n / (kms * 1e-3) / 1e9
        fl_ach = flops * n / (kms * 1e-3) / 1e12
        st = stats.cpu().numpy()
        subs = max(float(st[7]), 1.0) * 50
        ep = float(st[0])
        out = {
            "metric": "jaco_picking_env_steps_per_sec" if policy is None else "jaco_picking_policy_in_loop_env_steps_per_sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": (f"batched {args.task} roll-out: {n} envs per GPU (BASELINE config 3), reset distribution, actions U(-1,1)^7, "
                                    "auto-reset, 50 substeps x 1 ms per step; finger-object contact fraction in episode_stats, contact-rich mode in contact_rich")
                                   if policy is None else
                                   (f"HPC primitive-fusion policy roll-out in the loop: {n} envs per GPU (BASELINE config 5), composite policy (reaching 6-128^3-12, "
                                    "grasping 11-128^3-14, meta 17-256^3-3; random-initialised weights of the saved architecture, stochastic actions) "
                                    "forward in PyTorch fp32 + env step on the device, no host round trip"),
                       "baseline_config": args.config,
                       "actions": ("policy" if policy is not None else ("fresh i.i.d. U(-1,1)^7 batch per step, drawn on the device outside the timed region "
                                   "(like the CPU arm's per-step rng.uniform)" if not bank else
                                   f"bank of {bank} pre-generated batches, cycled (the rounds 1-2 protocol: periodic actions walk the arms into the singular posture)")),
                       "envs_per_gpu": n, "substeps_per_step": 50,
                       "episode_age": f"{args.age_steps} untimed roll-out steps after reset() before the {W} warm-up steps "
                                      "(step time grows with episode age, profiles/r01_step_time_trend.log)",
                       "l2": "flushed between timed iterations (256 MB fill outside the per-step CUDA-event pairs; "
                             f"state working set {116 * 4 * n / 1e6:.0f} MB)"},
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": n * NACT * 4,
                    "d2h_bytes_per_step": n * (NOBS * 4 + 4 + 4 + 1 + 1), "steps": K,
                    "protocol": "alternating with the device-timed steps on the same aged population (same step count, same L2 flush); host clock around jaco_step_host (H2D + step + D2H + stream sync)"},
            "gpu_launches": int(launches),
            # the binding roof is the FP32 pipe (instruction issue), not HBM -- roofline carries max(fp32 frac, hbm frac)
            "roofline": {"bound": "fp32", "achieved": fl_ach, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": fl_ach / fp32_peak if fp32_peak > 0 else None,
                         "traffic": prof.get("dram_bytes_per_launch") if n == 65536 else None,
                         "flops_per_env_step": flops,
                         "flops_source": (f"{prof['file']}: executed FADD + FMUL + 2 x FFMA (ncu hardware counters, aged workload)"
                                          if "fp32_flops_per_env_step" in prof else "SURVEY 8d a-priori estimate (no instrumented count committed)"),
                         "peak_source": "FMA loop measured in this run (jaco_fp32_peak_tflops)",
                         "note": "the kernel is bound by instruction delivery (DESIGN.md 5): both fractions are small by construction"},
            "roofline_hbm": {"bound": "hbm", "achieved": hbm_ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm_ach / peaks["hbm_gbs"],
                             "peak_source": which, "algorithmic_bytes_per_env_step": ALG_BYTES_PER_ENVSTEP},
            "kernel_ms": kms,
            "episode_stats": {"episodes": ep, "successes": float(st[1]), "timeouts": float(st[3]), "drops": float(st[4]),
                              "singular_exits": float(st[5]), "nan_resets": float(st[6]), "env_steps": float(st[7]),
                              "termination_causes": {"success": float(st[1]), "time_out": float(st[3]), "drop": float(st[4]),
                                                     "singular": float(st[5]), "other": ep - float(st[1] + st[3] + st[4] + st[5])},
                              "contact_fraction": float(st[13]) / subs, "stageB_fraction": float(st[11]) / subs, "pinv_fraction": float(st[12]) / subs,
                              "contact_cap_hits": float(st[9]), "iteration_cap_hits": float(st[10]),
                              "general_path_fallbacks": float(st[14]), "general_resets": float(st[15])},
            "per_rank": {"ms_per_step": [r[0] for r in per_rank], "stageB_fraction": [r[1] for r in per_rank],
                         "pinv_fraction": [r[2] for r in per_rank], "contact_fraction": [r[3] for r in per_rank], "episodes": [r[4] for r in per_rank],
                         "ms_per_step_min": [r[0] for r in per_rank_hw], "ms_per_step_max": [r[1] for r in per_rank_hw],
                         "sm_mhz": [r[2] for r in per_rank_hw], "throttle_reasons": [r[3] for r in per_rank_hw]},
            "contact_rich": contact_rich,
            "clocks": clk,
        }
        if not args.no_cpu_baseline and world == 1 and policy is None:
            out["cpu_baseline"] = cpu_baseline(args.cpu_seconds)
        print(json.dumps(out), file=_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
