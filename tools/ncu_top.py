"""Summarise an .ncu-rep: per kernel the instructions with most stall samples (SASS view)."""
import csv, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25; which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(out.splitlines()):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": [], "hdr": None}; blocks.append(cur); continue
    if cur is None: continue
    if cur["hdr"] is None: cur["hdr"] = row; continue
    cur["rows"].append(row)
b = blocks[which]
h = b["hdr"]; iS = h.index("# Samples"); iSrc = h.index("Source"); iEx = h.index("Instructions Executed")
tot = sum(int(r[iS] or 0) for r in b["rows"])
print(b["name"][:90], "total samples", tot, "n instr", len(b["rows"]))
idx = sorted(range(len(b["rows"])), key=lambda i: -int(b["rows"][i][iS] or 0))[:topn]
for i in sorted(idx):
    r = b["rows"][i]
    print(f"{i:5d} {int(r[iS]):6d} {100*int(r[iS])/max(tot,1):5.1f}%  ex={r[iEx]:>8}  {r[iSrc].strip()[:110]}")
