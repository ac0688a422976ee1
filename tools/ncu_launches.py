"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (developer tool)."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, agg = None, collections.OrderedDict()
for r in rows:
    if r and r[0] == "ID":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    d = dict(zip(hdr, r))
    a = agg.setdefault(d["Kernel Name"][:96], [0, 0.0])
    a[0] += 1
    a[1] += float(d["Metric Value"].replace(",", ""))
tot = sum(a[1] for a in agg.values())
print("per-kernel device time (ns, cold-cache serialised ncu replays; compare SHARES)")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-98s n=%3d total=%14.0f share=%.4f" % (k, n, t, t / tot))
