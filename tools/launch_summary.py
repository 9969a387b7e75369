"""Per-kernel summary of an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file ...)."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h, data = rows[hi], rows[hi + 1:]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.OrderedDict()
for r in data:
    if len(r) > vi:
        agg.setdefault(r[ki][:90], []).append(float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1e-3))
tot = sum(sum(v) for v in agg.values())
print(f"{len(data)} launches, {tot / 1e3:.2f} ms in total (cold-cache, serialised: compare shares)")
for k, v in agg.items():
    print(f"{len(v):4d} x  mean {sum(v) / len(v):9.1f} us  share {100 * sum(v) / tot:5.1f}%  {k}")
