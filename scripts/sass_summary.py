"""SASS instruction summary of the tcgen05 / TMEM / TMA kernels in code-nerf_b200/lib/libcnrf.so (CPU only: cuobjdump).
    python scripts/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "code-nerf_b200", "lib", "libcnrf.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ("UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTCBAR", "UTCATOMSWS", "SYNCS", "ELECT", "REDG", "ATOMG", "LDGSTS")
funcs, cur = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if cur and m:
        op = m.group(1).split(".")[0]
        funcs[cur]["n"] += 1
        if op in KEYS:
            funcs[cur][op] += 1
names = subprocess.run(["c++filt"], input="\n".join(funcs), capture_output=True, text=True).stdout.splitlines()
print("SASS instruction summary of the tcgen05 / TMEM / TMA kernels in code-nerf_b200/lib/libcnrf.so (cuobjdump -sass, sm_100a; build of this commit)")
print("UTCHMMA = tcgen05.mma kind::f16, LDTM/STTM = tcgen05.ld/st, UBLKCP = 1-D TMA bulk copy (cp.async.bulk), UTCBAR = tcgen05.commit,")
print("UTCATOMSWS = TMEM allocation, SYNCS = mbarrier ops, ELECT = elect.sync, REDG = red.global\n")
rows = [(n, c) for n, c in zip(names, funcs.values()) if c["UTCHMMA"] or c["UBLKCP"] or c["LDTM"]]
for n, c in sorted(rows, key=lambda r: -r[1]["n"]):
    print(n)
    print(f"    {c['n']} instructions: " + ", ".join(f"{k} {c[k]}" for k in KEYS if c[k]))
