"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv)."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]; kn = hdr.index("Kernel Name"); mv = hdr.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
  if len(r) <= mv: continue
  try: v = float(r[mv].replace(",", ""))
  except ValueError: continue
  a = agg.setdefault(r[kn], [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print(" ".join(sys.argv[2:]))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
  print("%-62s launches %4d  time %12.0f ns  share %.3f" % (k[:60], a[0], a[1], a[1] / tot))
