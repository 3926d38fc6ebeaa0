"""Full SASS listing of every kernel the training steps launch, one gzip per kernel under profiles/sass/ (the hex
encoding column is dropped; `cuobjdump -sass` of the built library is the source).  Conv-kernel template instantiations
that no engine launches are skipped (they differ only in unroll factors); pass --all to keep them.

    python scripts/dump_sass.py [--all]
"""
import gzip
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributedpytorch_b200", "_native", "libdpt_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
# igemm_conv_kernel<S, KSTEPS, VARIANT, CLUSTER> instantiations seen in the ncu captures of both models + the cluster ones
CONV_USED = {"4,1,0,1", "4,2,0,1", "2,2,0,1", "2,4,0,1", "4,4,0,1", "1,4,3,1", "2,4,3,1", "4,2,3,1", "1,4,1,1", "1,4,2,1",
             "2,2,2,1", "1,2,2,1", "1,4,1,2", "2,4,1,2", "1,4,1,4", "1,4,0,1", "1,2,0,1", "2,4,1,1"}
os.makedirs(OUT, exist_ok=True)
txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
demangle = {}
names = sorted(set(re.findall(r"Function : (\S+)", txt)))
dm = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
demangle = dict(zip(names, dm))
parts = re.split(r"(?=\t*Function : )", txt)
index = []
for part in parts:
    m = re.match(r"\s*Function : (\S+)", part)
    if not m:
        continue
    full = demangle.get(m.group(1), m.group(1))
    short = re.sub(r"^void ", "", full).split("(")[0].replace("dpt::", "")
    mm = re.search(r"igemm_conv_kernel<(\d+), (\d+), (\d+), (\d+)>", short.replace("(int)", ""))
    if mm and "--all" not in sys.argv and ",".join(mm.groups()) not in CONV_USED:
        continue
    fname = re.sub(r"[^A-Za-z0-9_]+", "_", short.replace("(int)", "")).strip("_")[:90] + ".sass.gz"
    body = re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/", "", part)
    with gzip.open(os.path.join(OUT, fname), "wt") as f:
        f.write(f"// {full}\n" + body)
    ops = re.findall(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", body)
    marks = {k: sum(1 for o in ops if o.startswith(k)) for k in ("UTCHMMA", "UTMALDG", "LDTM", "UTCBAR", "SYNCS", "LDGMC", "HMMA")}
    index.append((short.replace("(int)", ""), fname, len(ops), marks))
with open(os.path.join(OUT, "INDEX.md"), "w") as f:
    f.write("# SASS listings (cuobjdump -sass of libdpt_b200.so, sm_100a), one file per launched kernel\n\n")
    f.write("| kernel | file | instructions | UTCHMMA (tcgen05.mma) | UTMALDG (TMA) | LDTM (tcgen05.ld) | UTCBAR (commit) | SYNCS (mbarrier) | HMMA (legacy) |\n|---|---|---|---|---|---|---|---|---|\n")
    for short, fname, n, mk in sorted(index):
        f.write(f"| `{short}` | {fname} | {n} | {mk['UTCHMMA']} | {mk['UTMALDG']} | {mk['LDTM']} | {mk['UTCBAR']} | {mk['SYNCS']} | {mk['HMMA']} |\n")
print(f"{len(index)} kernels -> {OUT}")
