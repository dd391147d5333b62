#!/usr/bin/env python
"""Native-code evidence without a GPU: per-kernel counts of the Blackwell SASS mnemonics in lib/libftk.so
(cuobjdump -sass) and the ptxas register / spill / shared-memory table from the build logs.
usage: python tools/sass_summary.py <tag>   ->  profiles/<tag>_sass_summary.txt"""
import collections
import glob
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "kubernetes-cloud_b200")
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
MNEMONICS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "STTM", "UTCATOMSWS", "SYNCS", "UCGABAR",
             "MUFU.EX2", "HMMA", "LDGSTS", "ATOMG", "REDG", "STG.E.128", "STG.E.ENL2.256", "LDG.E.128"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


sass = subprocess.run(["cuobjdump", "-sass", os.path.join(PKG, "lib", "libftk.so")], capture_output=True, text=True).stdout
counts = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        counts[cur]["_total"] += 1
        for mn in MNEMONICS:
            if op.startswith(mn):
                counts[cur][mn] += 1
names = demangle(list(counts))
lines = ["# cuobjdump -sass kubernetes-cloud_b200/lib/libftk.so (sm_100a): instructions per kernel, Blackwell mnemonics",
         "# UTCHMMA = tcgen05.mma (bf16), UTCBAR = tcgen05.commit, UTMALDG/UTMASTG = TMA tensor load/store,",
         "# LDTM/STTM = tcgen05.ld/st (TMEM), SYNCS = mbarrier, HMMA = legacy mma.sync (must be 0 everywhere)", ""]
for fn, c in counts.items():
    pretty = re.sub(r"\(anonymous namespace\)::", "", names.get(fn, fn))
    pretty = pretty[: pretty.index("(")] if "(" in pretty and not pretty.startswith("void") else pretty.split("(CUtensorMap")[0]
    shown = "  ".join(f"{k}={v}" for k, v in c.items() if k != "_total" and v)
    lines.append(f"{pretty[:110]:110s} total={c['_total']:5d}  {shown}")
agg = collections.Counter()
for c in counts.values():
    agg.update({k: v for k, v in c.items() if k != "_total"})
lines += ["", "library totals: " + "  ".join(f"{k}={agg[k]}" for k in MNEMONICS if agg[k]), ""]
lines.append("# ptxas -v (registers / spills / static shared memory) from kubernetes-cloud_b200/build/*.ptxas.log")
for log in sorted(glob.glob(os.path.join(PKG, "build", "*.ptxas.log"))):
    txt = open(log).read()
    fns = re.findall(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers(.*)", txt)
    dm = demangle([f[0] for f in fns])
    for f in fns:
        pretty = re.sub(r"\(anonymous namespace\)::", "", dm.get(f[0], f[0])).split("(CUtensorMap")[0].split("(")[0]
        lines.append(f"{os.path.basename(log)[:-10]:10s} {pretty[:96]:96s} regs={f[4]:>3s} stack={f[1]:>3s} spill_st={f[2]} spill_ld={f[3]} {f[5].strip(', ')[:60]}")
out = os.path.join(ROOT, "profiles", f"{tag}_sass_summary.txt")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
