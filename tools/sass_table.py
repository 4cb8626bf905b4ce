"""profiles/sass_tcgen05.txt: per-kernel counts of the Blackwell-only SASS mnemonics in the shipped libnerf_b200.so
(UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk, UTCATOMSWS = TMEM alloc,
SYNCS = mbarrier ops; HMMA = mma.sync, here HMMA.1688.F32.TF32 of the EG3D decoder).  Usage: python tools/sass_table.py > profiles/sass_tcgen05.txt"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "nerf_siren_b200", "libnerf_b200.so")
MN = ["UTCHMMA", "HMMA", "LDTM", "UTCBAR", "UBLKCP", "UTCATOMSWS", "SYNCS", "MEMBAR", "REDG", "LDG", "STG"]
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
rows, cur, counts = [], None, None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        if cur:
            rows.append((cur, counts))
        cur, counts = m.group(1), dict.fromkeys(MN, 0)
        counts["instr"] = 0
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        counts["instr"] += 1
        for k in MN:
            if re.search(r"\b" + k + r"\b|\b" + k + r"\.", line):
                counts[k] += 1
if cur:
    rows.append((cur, counts))
print("# cuobjdump -sass nerf_siren_b200/libnerf_b200.so (sm_100a): occurrences per kernel")
print("%-86s %7s " % ("kernel", "instr") + " ".join("%10s" % k for k in MN))
for name, c in sorted(rows, key=lambda r: -r[1]["UTCHMMA"] * 100000 - r[1]["instr"]):
    d = demangle(name)
    d = re.sub(r"\(anonymous namespace\)::", "", d)
    d = d[:d.rfind("(")] if "(" in d else d
    d = d.replace("(bool)", "").replace("void ", "")[:86]
    print("%-86s %7d " % (d, c["instr"]) + " ".join("%10d" % c[k] for k in MN))
