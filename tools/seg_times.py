"""Per-segment launch times of the split pipeline from an ncu launch list (csv)."""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]; kn = hdr.index("Kernel Name"); mv = hdr.index("Metric Value")
SEG = ["act", "solve", "integrate+kin", "com+crb", "factor", "collide", "constraint", "vel1", "vel2"]
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
  if len(r) <= mv: continue
  try: v = float(r[mv].replace(",", ""))
  except ValueError: continue
  name = r[kn]
  mm = re.search(r"bp_seg_kernel<(\d+), *(\d+)>", name)
  key = "G%s %s" % (mm.group(1), SEG[int(mm.group(2))]) if mm else name[:40]
  a = agg.setdefault(key, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
for k, a in agg.items():
  print("%-28s launches %5d  mean %9.1f us  share %.3f" % (k, a[0], a[1] / a[0] / 1e3, a[1] / tot))
