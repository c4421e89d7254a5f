"""Summarise an `ncu --set full` report of the step kernel into the JSON bench.py reads (profiles/rNN_ncu_step_kernel.json).
usage: python scripts/ncu_summary.py report.ncu-rep n_envs out.json"""
import csv, io, json, subprocess, sys

rep, n_envs, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units, v = rows[0], rows[1], rows[-1]
col = {k: i for i, k in enumerate(h)}


def val(k, default=None):
    if k not in col:
        return default
    x = float(v[col[k]].replace(",", ""))
    u = units[col[k]].lower()
    scale = {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12, "byte": 1.0}.get(u, 1.0)
    return x * scale


fadd, fmul, ffma = (val(f"smsp__sass_thread_inst_executed_op_{k}_pred_on.sum", 0.0) for k in ("fadd", "fmul", "ffma"))
d = {
    "kernel": v[col["Kernel Name"]], "n_envs": n_envs,
    "duration_ms_under_ncu": val("gpu__time_duration.sum") / 1e6 if units[col["gpu__time_duration.sum"]] in ("ns", "nsecond") else val("gpu__time_duration.sum"),
    "registers_per_thread": val("launch__registers_per_thread"), "block_size": val("launch__block_size"), "grid_size": val("launch__grid_size"),
    "dram_bytes_per_launch": (val("dram__bytes_read.sum", 0) + val("dram__bytes_write.sum", 0)),
    "dram_bytes_read": val("dram__bytes_read.sum"), "dram_bytes_write": val("dram__bytes_write.sum"),
    "thread_inst_fadd": fadd, "thread_inst_fmul": fmul, "thread_inst_ffma": ffma,
    "fp32_flops_per_launch": fadd + fmul + 2 * ffma, "fp32_flops_per_env_step": (fadd + fmul + 2 * ffma) / n_envs,
    "warp_inst_executed": val("smsp__inst_executed.sum"), "threads_per_inst": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "issue_active_pct": val("smsp__issue_active.avg.pct"), "warps_active_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active"),
    "fma_pipe_pct": val("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    "local_load_inst": val("smsp__inst_executed_op_local_ld.sum"), "local_store_inst": val("smsp__inst_executed_op_local_st.sum"),
    "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
    "stalls_per_issue": {k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): float(v[i])
                         for k, i in col.items() if "issue_stalled" in k and k.endswith("_per_issue_active.ratio") and "not_issued" not in k and float(v[i] or 0) > 0.01},
}
json.dump(d, open(out, "w"), indent=1)
print(json.dumps(d, indent=1))
