"""Summarise .ncu-rep captures into a markdown table (key metrics per kernel launch)."""
import csv
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "time_us", 1e-3), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor_%", 1),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%", 1), ("dram__bytes_read.sum", "dram_rd_MB", 1),
        ("dram__bytes_write.sum", "dram_wr_MB", 1), ("lts__t_sector_hit_rate.pct", "l2_hit_%", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_%", 1), ("launch__registers_per_thread", "regs", 1),
        ("launch__grid_size", "grid", 1), ("launch__block_size", "block", 1),
        ("launch__shared_mem_per_block_dynamic", "dyn_smem_B", 1), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%", 1),
        ("smsp__inst_executed.sum", "inst", 1)]
print("| report | kernel | " + " | ".join(k[1] for k in KEYS) + " |")
print("|---|---|" + "---|" * len(KEYS))
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    if len(rows) < 3:
        continue
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        name = d.get("Kernel Name", "?").replace("(anonymous namespace)::", "").replace("<unnamed>::", "")[:60]
        vals = []
        for k, _, scale in KEYS:
            v = d.get(k, "")
            try:
                f = float(v.replace(",", ""))
                unit = u.get(k, "")
                if k.startswith("dram__bytes"):
                    f = f * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(unit, 1)
                elif k == "gpu__time_duration.sum":
                    f = f * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(unit, 1)
                vals.append(f"{f:.1f}" if abs(f) < 1e5 else f"{f:.3g}")
            except ValueError:
                vals.append(v)
        print(f"| {rep.split('/')[-1]} | {name} | " + " | ".join(vals) + " |")
