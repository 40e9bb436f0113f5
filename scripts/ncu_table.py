#!/usr/bin/env python
"""ncu report -> one markdown table, one row per kernel name (median over its launches).
usage: python scripts/ncu_table.py gpurun_out/prof_step.ncu-rep profiles/ncu_step_all_kernels.md "title" """
import collections, csv, re, statistics, subprocess, sys
rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
txt = open(rep).read() if rep.endswith(".csv") else \
    subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, units, rows = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
M = [("us", "gpu__time_duration.sum", 1e-3), ("dram rd MB", "dram__bytes_read.sum", None), ("dram wr MB", "dram__bytes_write.sum", None),
     ("dram %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1), ("L2 %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
     ("tensor %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 1),
     ("warps act %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1), ("issue %", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1),
     ("regs", "launch__registers_per_thread", 1), ("grid", "launch__grid_size", 1)]
STALLS = {"long_sb": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
          "math_thr": "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
          "barrier": "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
          "short_sb": "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
          "wait": "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
          "no_inst": "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
          "sleep": "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
          "lg_thr": "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"}


def num(r, name):
    if name not in col:
        return None
    try:
        v = float(r[col[name]].replace(",", ""))
    except ValueError:
        return None
    u = units[col[name]]
    if name.startswith("dram__bytes"):
        v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(u, 1)
    if name == "gpu__time_duration.sum":
        v *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(u, 1)
    return v


per = collections.OrderedDict()
for r in rows:
    name = re.sub(r"^(void )?(<unnamed>::)?", "", r[col["Kernel Name"]]).split("(")[0]
    per.setdefault(name, []).append(r)
with open(out, "w") as f:
    f.write(f"# {title}\n\n`ncu --clock-control none` with the SpeedOfLight, MemoryWorkloadAnalysis, WarpStateStats, SchedulerStats, LaunchStats and Occupancy sections (raw page of the report exported on the GPU box: {rep}, not tracked). One row per kernel: "
            f"median over its launches in the captured range, `n` = launches, `sum us` = their total. Serialised, cold-cache replay: "
            f"durations are not step time.  top stall = largest `smsp__average_warps_issue_stalled_*_per_issue_active` reason.\n\n")
    f.write("| kernel | n | sum us | " + " | ".join(m[0] for m in M) + " | top stalls |\n|---|---:|---:|" + "---:|" * len(M) + "---|\n")
    for name, rs in per.items():
        cells = []
        tot = sum(num(r, "gpu__time_duration.sum") or 0 for r in rs)
        for label, metric, _ in M:
            vals = [num(r, metric) for r in rs]
            vals = [v for v in vals if v is not None]
            cells.append(f"{statistics.median(vals):.1f}" if vals else "-")
        st = {}
        for k, metric in STALLS.items():
            vals = [num(r, metric) for r in rs]
            vals = [v for v in vals if v is not None]
            if vals:
                st[k] = statistics.median(vals)
        top = ", ".join(f"{k} {v:.2f}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        f.write(f"| `{name[:48]}` | {len(rs)} | {tot:.1f} | " + " | ".join(cells) + f" | {top} |\n")
print(open(out).read()[:3000])
