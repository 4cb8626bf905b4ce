"""Hot source lines of one kernel from an ncu report captured with --import-source on (library built with -lineinfo):
executed warp-instructions per CUDA source line.  Usage: python tools/ncu_hot_lines.py X.ncu-rep <kernel substring> [top]"""
import collections
import csv
import io
import subprocess
import sys

rep, want = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
per, text = collections.Counter(), {}
cur, hdr, ok = None, None, False
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r[0] == "Function Name":
        ok = want in r[1]
    elif r[0] == "Line No":
        hdr = r
    elif ok and hdr and r[0].isdigit():
        try:
            v = int(r[hdr.index("Instructions Executed")])
        except ValueError:
            v = 0
        per[(cur, int(r[0]))] += v
        text[(cur, int(r[0]))] = r[1]
tot = sum(per.values()) or 1
print(f"{want}: {tot} warp-instructions")
for (f, l), v in per.most_common(top):
    print(f"{v / tot * 100:5.1f}%  {f}:{l}  {text[(f, l)].strip()[:120]}")
