"""Summarise an .ncu-rep for profiles/: per kernel launch the counters BASELINE.json asks for (L2 hit rate, achieved L2 / HBM GB/s,
warp execution efficiency, issue-slot utilisation) plus occupancy, stall reasons and DRAM traffic.
usage: ncu_summary.py file.ncu-rep [--json traffic.json [--config cfg4] [--source "text"]]"""
import csv, json, subprocess, sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, units = rows[0], rows[1]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0,
         "byte/second": 1, "Kbyte/second": 1e3, "Mbyte/second": 1e6, "Gbyte/second": 1e9, "Tbyte/second": 1e12}


def val(r, k):
    if k not in h:
        return None
    i = h.index(k)
    try:
        v = float(r[i].replace(",", ""))
    except ValueError:
        return None
    return v * SCALE.get(units[i], 1)


per = {}
for r in rows[2:]:
    name = r[h.index("Kernel Name")].split("(")[0].replace("void ", "")
    t = val(r, "gpu__time_duration.sum")
    d = dict(
        ms=t * 1e3, warp_inst=val(r, "smsp__inst_executed.sum"),
        issue_active_pct=val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        occupancy_pct=val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        active_lanes=val(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
        l1_hit_pct=val(r, "l1tex__t_sector_hit_rate.pct"), l2_hit_pct=val(r, "lts__t_sector_hit_rate.pct"),
        l2_gbs=(val(r, "lts__t_sectors.sum") or 0) * 32 / t / 1e9, l2_throughput_pct=val(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
        l1_gbs=(val(r, "l1tex__t_bytes_pipe_lsu_mem_global_op_ldgsts_cache_access.sum") or 0) / t / 1e9,
        dram_read_bytes=val(r, "dram__bytes_read.sum"), dram_write_bytes=val(r, "dram__bytes_write.sum"),
        dram_gbs=((val(r, "dram__bytes_read.sum") or 0) + (val(r, "dram__bytes_write.sum") or 0)) / t / 1e9,
        regs=val(r, "launch__registers_per_thread"),
        stall_long_scoreboard=val(r, "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
        stall_wait=val(r, "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"),
        stall_not_selected=val(r, "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"),
        stall_math_throttle=val(r, "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"),
        stall_branch=val(r, "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio"),
        stall_short_scoreboard=val(r, "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"),
        stall_barrier=val(r, "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"),
        stall_dispatch=val(r, "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio"),
        stall_no_inst=val(r, "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"),
        pipe_alu_pct=val(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        pipe_fma_pct=val(r, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        pipe_fmaheavy_pct=val(r, "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active"),
        pipe_lsu_pct=val(r, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
        pipe_xu_pct=val(r, "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
        pred_on_lanes=val(r, "smsp__thread_inst_executed_pred_on_per_inst_executed.ratio"))
    per.setdefault(name, []).append(d)
    print(name)
    print("   " + "  ".join(f"{k}={v:.4g}" if isinstance(v, float) else f"{k}={v}" for k, v in d.items()))
if "--json" in sys.argv:
    # per-launch DRAM traffic of each kernel (mean over the captured launches; reflection waves: the first = largest wave)
    cfgname = sys.argv[sys.argv.index("--config") + 1] if "--config" in sys.argv else "cfg4"
    jpath = sys.argv[sys.argv.index("--json") + 1]
    try:
        tr = json.load(open(jpath))
    except Exception:
        tr = {}
    tr["source"] = sys.argv[sys.argv.index("--source") + 1] if "--source" in sys.argv else "ncu --set full capture: " + sys.argv[1]
    for name, ds in per.items():
        key = name.split("<")[0]
        big = ds if "bounce" not in key else ds[:1]
        tr[cfgname + ":" + key + "_dram_bytes_per_launch"] = sum((d["dram_read_bytes"] or 0) + (d["dram_write_bytes"] or 0) for d in big) / len(big)
    json.dump(tr, open(jpath, "w"), indent=1)
