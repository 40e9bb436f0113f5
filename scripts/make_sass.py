#!/usr/bin/env python
"""Full SASS listing of every kernel in crackfl/_lib/libcrackfl.so, one file per kernel under profiles/sass/ (instruction
text only: addresses and encodings stripped so that a rebuild only changes what the compiler changed), plus
profiles/sass_mnemonics.md: per kernel, the counts of the mnemonics that show which hardware path it uses.

  UTC*MMA = tcgen05.mma   LDTM/STTM = tcgen05.ld/st   UTMALDG/UTMASTG = TMA tensor load/store   UBLKCP = cp.async.bulk
  SYNCS = mbarrier   LDGMC/MULTIMEM = multimem.ld_reduce / multimem.st (NVLS)   RED/ATOM = global reductions
  *.SYS = system-scope (cross-GPU) loads/stores   FFMA2/FADD2 = packed fp32
Run on the CPU box after a build: python scripts/make_sass.py"""
import collections, os, re, shutil, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "crackfl", "_lib", "libcrackfl.so")
OUT = os.path.join(ROOT, "profiles", "sass")
PAT = re.compile(r"\b(UTC[A-Z]*MMA|UTMALDG|UTMASTG|UTMAPF|UBLKCP|UBLKRED|LDTM|STTM|UTCBAR|UTCCP|SYNCS|HMMA|LDGSTS|LDGMC|RED|REDG|ATOM|ATOMG|ATOMS|"
                 r"MULTIMEM|FFMA2|FADD2|FMUL2|MUFU|CCTL|ERRBAR|MEMBAR|ACQBULK|ARRIVES|ELECT)\b")


def short(name):
    n = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
    n = re.sub(r"\(anonymous namespace\)::", "", n)
    n = re.sub(r"\(.*", "", n)
    n = re.sub(r"^void ", "", n)
    return n


def main():
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    shutil.rmtree(OUT, ignore_errors=True)
    os.makedirs(OUT, exist_ok=True)
    funcs, cur = collections.OrderedDict(), None
    for ln in txt.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(.*?);\s*/\*", ln)
        if m:
            funcs[cur].append(m.group(1).strip())
        elif re.match(r"\s*\.L_x_\d+:", ln):
            funcs[cur].append(ln.strip())
    rows = []
    for mangled, ins in funcs.items():
        name = short(mangled)
        fn = re.sub(r"[^A-Za-z0-9_.]+", "_", name).strip("_")[:80] + ".sass"
        cnt = collections.Counter()
        sys_ops = 0
        for i in ins:
            for mm in PAT.findall(i):
                cnt[mm] += 1
            if re.search(r"\b(LDG|STG|LD|ST)\S*\.SYS\b|\.STRONG\.SYS", i):
                sys_ops += 1
        if sys_ops:
            cnt["*.SYS ld/st"] = sys_ops
        with open(os.path.join(OUT, fn), "w") as f:
            f.write(f"// {name}\n// {mangled}\n// cuobjdump -sass crackfl/_lib/libcrackfl.so (sm_100a), {len(ins)} lines, addresses/encodings stripped\n")
            f.write("\n".join(ins) + "\n")
        rows.append((name, fn, len(ins), cnt))
    with open(os.path.join(ROOT, "profiles", "sass_mnemonics.md"), "w") as f:
        f.write("# SASS evidence, every kernel of libcrackfl.so (sm_100a)\n\nFull listings: `profiles/sass/<kernel>.sass` "
                "(`python scripts/make_sass.py`).  UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA tensor "
                "load/store, UBLKCP = cp.async.bulk, SYNCS = mbarrier, LDGMC / MULTIMEM = multimem.ld_reduce / multimem.st "
                "(NVLS), RED/ATOM = global reductions, `*.SYS` = system-scope (peer) accesses, FFMA2 = packed fp32.\n\n"
                "| kernel | SASS lines | mnemonics | listing |\n|---|---:|---|---|\n")
        for name, fn, n, cnt in sorted(rows):
            f.write(f"| `{name[:64]}` | {n} | {', '.join(f'{m} x{c}' for m, c in sorted(cnt.items())) or '-'} | `sass/{fn}` |\n")
    print(f"{len(rows)} kernels -> {OUT}")


if __name__ == "__main__":
    main()
