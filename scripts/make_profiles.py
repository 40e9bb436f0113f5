#!/usr/bin/env python
"""Condense gpurun_out/ artefacts (ncu launch lists, ncu reports, sweeps, bench lines) into tracked summaries under
profiles/.  Run on the CPU box after a gpurun call:  python scripts/make_profiles.py"""
import collections, csv, glob, json, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
GO = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)


def launches(path, tag):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    for row in csv.DictReader(lines):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((row["Kernel Name"], float(row["Metric Value"].replace(",", "")), row.get("Grid Size")))
    idx = [i for i, r in enumerate(rows) if "adam_k" in r[0]]
    if len(idx) < 2:
        return
    step = rows[idx[-2] + 1: idx[-1] + 1]
    tot = sum(r[1] for r in step)
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in step:
        k = re.sub(r"\(.*", "", r[0]).split("::")[-1]
        agg[k][0] += 1
        agg[k][1] += r[1]
    with open(os.path.join(OUT, f"step_kernels_{tag}.md"), "w") as f:
        f.write(f"# Train-step kernel list ({tag})\n\nSource: `ncu --metrics gpu__time_duration.sum --clock-control none` on "
                f"`bench.py --steps 3 --warmup 3 --no-e2e` (U-Net 256x256, B=16, bf16, 1 B200). Serialised, cold-cache: "
                f"compare shares, not absolutes.\n\n{len(step)} kernels in the last full step, {tot/1000:.1f} us summed.\n\n"
                f"| kernel | launches | us | share |\n|---|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k}` | {v[0]} | {v[1]/1000:.1f} | {100*v[1]/tot:.1f}% |\n")


def ncu_raw(rep, tag, title):
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__grid_size", "lts__t_sectors_srcunit_tex_op_read.sum",
            "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]
    try:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    except FileNotFoundError:
        return
    rows = list(csv.reader(txt.splitlines()))
    if len(rows) < 3:
        return
    hdr, units = rows[0], rows[1]
    with open(os.path.join(OUT, f"ncu_{tag}.md"), "w") as f:
        f.write(f"# {title}\n\n`ncu --set full --clock-control none --import-source on` (report: gpurun_out/{os.path.basename(rep)}, "
                f"not tracked).\n\n")
        for r in rows[2:]:
            name = r[hdr.index("Kernel Name")].split("(")[0].split("::")[-1]
            f.write(f"## {name}  (launch id {r[hdr.index('ID')]})\n\n| metric | value | unit |\n|---|---:|---|\n")
            for w in want:
                if w in hdr:
                    i = hdr.index(w)
                    f.write(f"| `{w}` | {r[i]} | {units[i]} |\n")
            f.write("\n")


def sass():
    so = os.path.join(ROOT, "crackfl", "_lib", "libcrackfl.so")
    if not os.path.exists(so):
        return
    txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    cur, per = None, collections.defaultdict(collections.Counter)
    pat = re.compile(r"\b(UTC[A-Z]*MMA|UTMALDG|UTMASTG|UBLKCP|LDTM|STTM|UTCBAR|UTCCP|SYNCS|HMMA|LDGSTS|RED|ATOM[GS]?|MULTIMEM|UTMAPF)\b")
    for ln in txt.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            continue
        if cur:
            for mm in pat.findall(ln):
                per[cur][mm] += 1
    with open(os.path.join(OUT, "sass_mnemonics.md"), "w") as f:
        f.write("# SASS evidence (cuobjdump -sass crackfl/_lib/libcrackfl.so, sm_100a)\n\nUTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, "
                "UTMALDG/UTMASTG = TMA load/store, SYNCS = mbarrier, MULTIMEM = multimem.st (NVLS).\n\n| kernel | mnemonics |\n|---|---|\n")
        for k, c in sorted(per.items()):
            if any(x.startswith(("UTC", "UTMA", "LDTM", "MULTIMEM")) for x in c):
                short = re.sub(r"^_ZN\d+_GLOBAL__N__[0-9a-f_]+", "", k)
                f.write(f"| `{short[:70]}` | {', '.join(f'{m} x{n}' for m, n in sorted(c.items()))} |\n")


def copy_json(pattern, dst):
    for fn in glob.glob(os.path.join(GO, pattern)):
        try:
            d = json.load(open(fn))
        except Exception:
            try:
                d = [json.loads(l) for l in open(fn) if l.strip().startswith("{")]
            except Exception:
                continue
        json.dump(d, open(os.path.join(OUT, dst or os.path.basename(fn)), "w"), indent=1)


if __name__ == "__main__":
    tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
    for fn in sorted(glob.glob(os.path.join(GO, "launches*.csv"))):
        launches(fn, os.path.basename(fn)[:-4].replace("launches", "v") or "v")
    for rep, t, title in (("prof_conv.ncu-rep", "conv3x3_d3ct1", "tcgen05 conv fwd/dgrad + wgrad, 3x3 64->32 @128^2 B=16"),
                          ("prof_pw.ncu-rep", "pointwise_e0pw2", "tcgen05 conv fwd/dgrad + wgrad, 1x1 64->64 @128^2 B=16"),
                          ("prof_dw.ncu-rep", "dw_stencil_gen1", "depthwise 3x3 (first TMA version: fully unrolled, activation inlined), 64 ch @128^2 B=16"),
                          ("prof_dw2.ncu-rep", "dw_gen2", "depthwise 3x3 fwd / dgrad / wgrad (pre-pass + FFMA2 version), 64 ch @128^2 B=16"),
                          ("prof_d3.ncu-rep", "conv3x3_d3ct1_halo_lane0", "halo conv 3x3 64->32 @128^2 before the warp-uniform issue path (MMA thread under if(lane==0))"),
                          ("prof_small.ncu-rep", "small_gemm_d1res", "1x1 256->128 @16^2 B=16: fixed cost of a small tcgen05 GEMM launch")):
        p = os.path.join(GO, rep)
        if os.path.exists(p):
            ncu_raw(p, t, title)
    sass()
    for pat in ("agg_sweep_*.json", "conv_bench*.json", "bench_*.json", "dw_bench*.json"):
        copy_json(pat, None)
    print(sorted(os.listdir(OUT)))
