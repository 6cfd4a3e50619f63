"""Summarise an ncu launch list (--csv, gpu__time_duration.sum) into per-kernel totals and shares."""
import csv, sys, collections, re, json
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
H = rows[hdr]
kn, mv, mu = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[hdr + 1:]:
    try:
        v = float(r[mv].replace(",", ""))
    except ValueError:
        continue
    u = r[mu]
    v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(u, 1.0)
    name = re.sub(r"\(.*", "", r[kn])
    a = tot.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
s = sum(a[1] for a in tot.values())
out = [{"kernel": k, "launches": a[0], "total_us": round(a[1], 2), "avg_us": round(a[1] / a[0], 3), "share": round(a[1] / s, 4)} for k, a in
       sorted(tot.items(), key=lambda kv: -kv[1][1])]
print(json.dumps({"total_us": round(s, 1), "kernels": out}, indent=1))
