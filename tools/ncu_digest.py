"""Digest one .ncu-rep (ncu --set full --import-source on) into small text files that fit the
gpurun_out budget: a metric table of every captured launch and, for the slowest launch of each
distinct kernel, the SASS/source lines with the most stall samples.

    python tools/ncu_digest.py <report.ncu-rep> <out_dir> [top_n_lines]
    python tools/ncu_digest.py <prefix>  <out_dir> [top_n_lines]   # reads <prefix>_raw.csv.gz / <prefix>_source.csv.gz
                                                                    # (ncu -i rep --page raw|source --csv | gzip)
"""
import gzip
import csv
import subprocess
import sys
from pathlib import Path

rep, out_dir = sys.argv[1], Path(sys.argv[2])
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 18
out_dir.mkdir(parents=True, exist_ok=True)

KEYS = [("gpu__time_duration.sum", "time_us"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor_%"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%"), ("dram__bytes_read.sum", "dram_rd_MB"),
        ("dram__bytes_write.sum", "dram_wr_MB"), ("lts__t_sector_hit_rate.pct", "l2_hit_%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_pipe_%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_%"), ("launch__registers_per_thread", "regs"),
        ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__cluster_size", "cluster"),
        ("launch__shared_mem_per_block_dynamic", "dyn_smem_KB"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%"),
        ("smsp__inst_executed.sum", "inst")]


def clean(name):
    return name.replace("(anonymous namespace)::", "").replace("<unnamed>::", "")




def page(which):
    if rep.endswith(".ncu-rep"):
        return subprocess.run(["ncu", "-i", rep, "--page", which, "--csv"], capture_output=True, text=True).stdout
    with gzip.open(f"{rep}_{which}.csv.gz", "rt") as fh:
        return fh.read()


raw = page("raw")
rows = list(csv.reader(raw.splitlines()))
hdr, units, body = rows[0], rows[1], rows[2:]
launches = []
with open(out_dir / "ncu_launch_table.md", "w") as f:
    f.write("| # | kernel | " + " | ".join(k[1] for k in KEYS) + " |\n|---|---|" + "---|" * len(KEYS) + "\n")
    for i, r in enumerate(body):
        d, u = dict(zip(hdr, r)), dict(zip(hdr, units))
        name = clean(d.get("Kernel Name", "?"))
        vals = []
        t_us = 0.0
        for k, label in KEYS:
            v = d.get(k, "")
            try:
                x = float(v.replace(",", ""))
                unit = u.get(k, "")
                if k.startswith("dram__bytes"):
                    x *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(unit, 1)
                elif k == "gpu__time_duration.sum":
                    x *= {"ns": 1e-3, "us": 1, "ms": 1e3}.get(unit, 1)
                    t_us = x
                elif label == "dyn_smem_KB":
                    x *= {"byte": 1e-3, "Kbyte": 1, "Mbyte": 1e3}.get(unit, 1)
                vals.append(f"{x:.1f}" if abs(x) < 1e5 else f"{x:.3g}")
            except ValueError:
                vals.append(v)
        launches.append((i, name, t_us))
        f.write(f"| {i} | {name[:70]} | " + " | ".join(vals) + " |\n")

# slowest launch of each distinct kernel
best = {}
for i, name, t in launches:
    key = name.split("(")[0]
    if key not in best or t > best[key][1]:
        best[key] = (i, t)

src = page("source")
blocks, cur = [], None
for row in csv.reader(src.splitlines()):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": [], "hdr": None}
        blocks.append(cur)
        continue
    if cur is None:
        continue
    if cur["hdr"] is None:
        cur["hdr"] = row
        continue
    cur["rows"].append(row)

# the source page may print a launch more than once (identical SASS views): drop exact repeats so
# block k belongs to launch k
dedup = []
for b in blocks:
    if dedup and dedup[-1]["name"] == b["name"] and dedup[-1]["rows"] == b["rows"]:
        continue
    dedup.append(b)
blocks = dedup
stride = 1
if len(blocks) != len(launches):
    print(f"warning: {len(blocks)} source blocks for {len(launches)} launches; stall listings may be misattributed")
with open(out_dir / "ncu_top_stalls.txt", "w") as f:
    for key, (i, t) in sorted(best.items(), key=lambda kv: -kv[1][1]):
        if i * stride >= len(blocks) or t < 5.0:
            continue
        b = blocks[i * stride]
        h = b["hdr"]
        try:
            iS, iSrc, iEx = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
        except ValueError:
            continue
        tot = sum(int(r[iS] or 0) for r in b["rows"])
        f.write(f"\n=== launch {i}: {clean(b['name'])[:110]}  ({t:.1f} us under ncu; {tot} stall samples, {len(b['rows'])} SASS lines)\n")
        idx = sorted(range(len(b["rows"])), key=lambda j: -int(b["rows"][j][iS] or 0))[:topn]
        for j in sorted(idx):
            r = b["rows"][j]
            f.write(f"{j:5d} {int(r[iS] or 0):6d} {100 * int(r[iS] or 0) / max(tot, 1):5.1f}%  ex={r[iEx]:>9}  {r[iSrc].strip()[:120]}\n")
print("launches:", len(launches), "distinct kernels:", len(best))
