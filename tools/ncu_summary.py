"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel mean and share."""
import collections, csv, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
agg = collections.OrderedDict()
for r in rows[1 + skip:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    name = r[ki].split("(")[0].replace("sigsvgd::<unnamed>::", "").replace("void ", "")
    agg.setdefault(name, []).append(v)
tot = sum(sum(v) for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k[:60]:60s} n={len(v):4d} mean_us={sum(v) / len(v) / 1000:9.1f} share={sum(v) / tot * 100:5.1f}%")
print(f"total_us={tot / 1000:.1f}")
