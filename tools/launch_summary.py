"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(',', ''))
    v = v / 1e3 if r[ui] == 'ns' else (v * 1e3 if r[ui] == 'ms' else v)
    k = r[ki][:70]; agg[k][0] += 1; agg[k][1] += v
tot = sum(t for _, t in agg.values())
print(f"total {tot/1e3:.3f} ms over {sum(c for c,_ in agg.values())} launches")
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{t/1e3:10.3f} ms {100*t/tot:5.1f}% {c:6d} launches  avg {t/c:9.1f} us  {k}")
