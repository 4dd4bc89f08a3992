"""Write profiles/<name>.md from an ncu --set full report: headline metrics of the profiled kernel plus the
per-source-line / per-barrier-segment breakdown of tools/ncu_lines.py.
Usage: python tools/ncu_summary.py REPORT.ncu-rep NAME 'kernel substring' 'command that was profiled' [translation unit]"""
import csv, subprocess, sys, os, json
rep, name, ksub, cmd = sys.argv[1:5]
obj = sys.argv[5] if len(sys.argv) > 5 else "pose_kernels"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, val = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio"]
kn = val[hdr.index("Kernel Name")]
lines = [f"# ncu --set full: {name}", "", f"Command profiled (after it exited 0 without ncu): `{cmd}`", "",
         f"Kernel: `{kn}`  (one launch, `--clock-control none`)", "", "| metric | value | unit |", "|---|---|---|"]
got = {}
for w in want:
    if w in hdr:
        i = hdr.index(w); got[w] = val[i]
        lines.append(f"| {w} | {val[i]} | {units[i]} |")
def num(x): return float(x.replace(",", ""))
def to_bytes(w):
    v = num(got[w]); u = units[hdr.index(w)]
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
traffic = to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum")
lines += ["", f"DRAM traffic of this launch: {traffic:.0f} bytes (read + write).", "",
          "## Hot source lines and barrier segments (tools/ncu_lines.py)", "", "```"]
lines += subprocess.run([sys.executable, os.path.join(root, "tools/ncu_lines.py"), rep, ksub, "1.0", obj], capture_output=True, text=True).stdout.splitlines()
lines += ["```", ""]
open(os.path.join(root, "profiles", name + ".md"), "w").write("\n".join(lines))
print(json.dumps({"traffic_bytes_per_launch": traffic, "kernel_ms": got.get("gpu__time_duration.sum")}))
