"""Aggregate an ncu launch list (gpu__time_duration.sum per launch) by kernel; second half only
(= the second of two identical steps) when --half is given."""
import collections, csv, sys
path = sys.argv[1]; half = "--half" in sys.argv
lines = [l for l in open(path) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
if half: rows = rows[len(rows) // 2:]
agg = collections.defaultdict(lambda: [0, 0.0]); tot = 0.0
for r in rows:
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    v = v / 1000 if u == "ns" else (v * 1000 if u == "ms" else v)
    n = r["Kernel Name"].replace("<unnamed>::", "").replace("void ", "")[:64]
    agg[n][0] += 1; agg[n][1] += v; tot += v
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
    print(f"{t:9.1f} us {100*t/tot:5.1f}%  n={n:4d}  avg={t/n:7.1f}  {k}")
print(f"total {tot:.1f} us over {len(rows)} launches")
