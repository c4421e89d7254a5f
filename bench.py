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
# algorithmic FLOPs per env-step (FMA = 2): instrumented count of the fp32 path on the picking roll-out
# distribution, see DESIGN.md "flop count" (tests/harness counts them); a-priori SURVEY estimate until measured
ALG_FLOPS_PER_ENVSTEP = float(os.environ.get("JACO_FLOPS_PER_ENVSTEP", 1.9e6))


def _ncu_traffic():
    """dram bytes (read + write) per launch of the step kernel from the committed ncu --set full summary, or None."""
    p = os.path.join(REPO, "profiles", "r01_ncu_step_kernel.json")
    try:
        return json.load(open(p)).get("dram_bytes_per_launch")
    except Exception:
        return None


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
    wid, task, seconds, seed = args
    import numpy as np
    from oracle.oracle import OracleEnv
    env = OracleEnv(task=task, seed=seed, env_id=wid)
    env.reset()
    rng = np.random.default_rng(wid)
    for _ in range(5):
        env.step(rng.uniform(-1, 1, 7))
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        _, _, d, _, _ = env.step(rng.uniform(-1, 1, 7))
        n += 1
        if d:
            env.reset()
    return n, time.perf_counter() - t0


def cpu_baseline(seconds, task="picking", workers=None):
    """Oracle (CPU restatement of the reference path), one env per process, all host cores."""
    import multiprocessing as mp
    from oracle import oracle as _o
    _o.build()
    workers = workers or os.cpu_count() or 1
    with mp.get_context("spawn").Pool(workers) as pool:
        res = pool.map(_oracle_worker, [(w, task, seconds, 0) for w in range(workers)])
    rate = sum(n / t for n, t in res)
    return {"value": rate, "unit": "env-steps/s", "cores": workers, "kind": "port",
            "sample": f"{workers} processes x 1 env, picking roll-out (reset distribution, U(-1,1)^7 actions, auto-reset), "
                      f"{seconds:.0f} s wall each; oracle/jaco_oracle.c float64 (CPU restatement; mujoco-py unavailable on this box)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step_s = max(2.0, min(10.0, 60.0 / max(1, args.steps + args.warmup)))
    vals = []
    for i in range(args.warmup + args.steps):
        cb = cpu_baseline(per_step_s)
        if i >= args.warmup:
            vals.append(cb["value"])
    v = sum(vals) / len(vals)
    cb["value"] = v
    print(file=_OUT, flush=True, *[json.dumps({
        "impl": "reference", "metric": "jaco_picking_env_steps_per_sec", "value": v, "unit": "env-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * per_step_s,
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
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    from mujoco_jaco_b200.shard import max_over_ranks, reduce_stats, shard_range

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries exactly one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    n, K, W = args.envs, args.steps, max(3, args.warmup)
    env_offset, _ = shard_range(rank, world, n)   # independent shard of the global env range: no data-path collective
    env = JacoVecEnv(n, task=args.task, seed=0, device=local, env_offset=env_offset, auto_reset=True)
    env.reset()
    gen = torch.Generator(device=dev); gen.manual_seed(1234 + rank)
    nact = K + W
    actions = torch.rand(min(nact, 32), n, NACT, device=dev, generator=gen) * 2 - 1   # resident in HBM before timing
    # de-synchronise episode phases a little so the timed region sees the steady-state mix of contact states
    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for i in range(args.age_steps):
        env.step(actions[i % actions.shape[0]])
    for i in range(W):
        env.step(actions[i % actions.shape[0]])
    env.stats(clear=True)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = env.launch_count()
    # L2 is flushed between timed iterations (a 256 MB fill, outside the per-step event pairs): the state working set
    # (30 MB) would otherwise stay resident in the 126 MB L2 from one step to the next
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    barrier()
    for i in range(K):
        flush.fill_(i & 255)
        ev0[i].record()
        env.step(actions[(W + i) % actions.shape[0]])
        ev1[i].record()
    barrier()
    kernel_ms = [ev0[i].elapsed_time(ev1[i]) for i in range(K)]
    total_ms = sum(kernel_ms)
    launches = env.launch_count() - launches0
    stats = env.stats(clear=True).clone()
    total_ms = max_over_ranks(total_ms, device=dev)   # device time, max over ranks
    reduce_stats(stats)                                # the only collective of the path: episode statistics (NCCL)
    value = world * n * K / (total_ms * 1e-3)

    # ---- e2e through the host-buffer C-ABI call (pinned host buffers, copies inside the timed region)
    Ke = max(3, min(K, 10))
    a_hosts = [torch.empty(n, NACT).uniform_(-1, 1).pin_memory() for _ in range(8)]   # fresh actions every step
    outs = (torch.empty(n, NOBS).pin_memory(), torch.empty(n).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory(),
            torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n).pin_memory())
    outs_np = tuple(o.numpy() for o in outs)
    a_nps = [a.numpy() for a in a_hosts]
    env.step_host(a_nps[-1], outs_np)
    barrier()
    t0 = time.perf_counter()
    for i in range(Ke):
        env.step_host(a_nps[i % len(a_nps)], outs_np)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    e2e_value = world * n * Ke / max_over_ranks(e2e_s, device=dev)
    if sampler:
        sampler.stop_flag = True
        sampler.join(timeout=2)

    if rank == 0:
        peaks, which = _peaks()
        L = _lib.load()
        fp32_peak = float(L.jaco_fp32_peak_tflops(local, 5))
        kms = sum(kernel_ms) / len(kernel_ms)
        hbm_ach = ALG_BYTES_PER_ENVSTEP * n / (kms * 1e-3) / 1e9
        fl_ach = ALG_FLOPS_PER_ENVSTEP * n / (kms * 1e-3) / 1e12
        st = stats.cpu().numpy()
        out = {
            "metric": "jaco_picking_env_steps_per_sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"batched {args.task} with finger-object contacts: {n} envs per GPU (BASELINE config 3), "
                                   "reset distribution, actions U(-1,1)^7, auto-reset, 50 substeps x 1 ms per step",
                       "envs_per_gpu": n, "substeps_per_step": 50,
                       "episode_age": f"{args.age_steps} untimed roll-out steps after reset() before the {W} warm-up steps "
                                      "(step time grows with episode age, profiles/r01_step_time_trend.log)",
                       "l2": "flushed between timed iterations (256 MB fill outside the per-step CUDA-event pairs; "
                             f"state working set {116 * 4 * n / 1e6:.0f} MB)"},
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": n * NACT * 4,
                    "d2h_bytes_per_step": n * (NOBS * 4 + 4 + 4 + 1 + 1), "steps": Ke},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": hbm_ach, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": hbm_ach / peaks["hbm_gbs"], "traffic": _ncu_traffic() if n == 65536 else None, "peak_source": which,
                         "note": "HBM is not the limiter of this kernel (50 substeps fused per state round trip); see fp32"},
            "roofline_fp32": {"bound": "fp32", "achieved": fl_ach, "peak": fp32_peak, "unit": "TFLOP/s",
                              "frac": fl_ach / fp32_peak if fp32_peak > 0 else None,
                              "flops_per_env_step": ALG_FLOPS_PER_ENVSTEP, "peak_source": "measured FMA loop (jaco_fp32_peak_tflops)"},
            "kernel_ms": kms,
            "episode_stats": {"episodes": float(st[0]), "successes": float(st[1]), "timeouts": float(st[3]),
                              "nan_resets": float(st[6]), "env_steps": float(st[7])},
            "clocks": sampler.summary() if sampler else None,
        }
        if not args.no_cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_baseline(args.cpu_seconds)
        print(json.dumps(out), file=_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
