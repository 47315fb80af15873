"""Aggregates an ncu source page by device function (needs the matching .o for symbols).
usage: ncu_by_function.py report.ncu-rep obj.o kernel_symbol_substring"""
import csv, subprocess, sys

def symbols(obj, kern):
  out = subprocess.run(["cuobjdump", "-elf", obj], capture_output=True, text=True).stdout
  syms = []
  for l in out.splitlines():
    p = l.split()
    if len(p) >= 7 and p[-1].startswith("$") and kern in p[-1]:
      try:
        syms.append((int(p[1], 16), int(p[2], 16), p[-1].split("$")[-1]))
      except ValueError:
        pass
  return sorted(syms)

def main(rep, obj, kern):
  syms = symbols(obj, kern)
  out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
  rows = list(csv.reader(out.splitlines()))
  hdr = rows[1]
  ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
  ith = hdr.index("Thread Instructions Executed")
  cols = {k: hdr.index(k) for k in ("stall_no_inst", "stall_long_sb", "stall_wait", "stall_short_sb", "stall_branch_resolving")}
  base = int(rows[2][ia], 16)
  agg = {}
  for r in rows[2:]:
    try:
      off = int(r[ia], 16) - base
    except (ValueError, IndexError):
      continue
    name = "<kernel>"
    for o, sz, n in syms:
      if o <= off < o + sz:
        name = n
        break
    a = agg.setdefault(name, dict(inst=0, thr=0, samp=0, **{k: 0 for k in cols}))
    a["inst"] += int(r[ii]); a["samp"] += int(r[isamp]); a["thr"] += int(r[ith])
    for k, i in cols.items():
      a[k] += int(r[i])
  tot_i = sum(a["inst"] for a in agg.values()); tot_s = sum(a["samp"] for a in agg.values())
  print("%-46s %7s %7s %6s | %s" % ("function", "inst%", "samp%", "lanes", " ".join(k[6:14] for k in cols)))
  for n, a in sorted(agg.items(), key=lambda kv: -kv[1]["samp"]):
    if a["samp"] == 0 and a["inst"] == 0:
      continue
    print("%-46s %7.2f %7.2f %6.1f | %s" % (n[:46], 100.0 * a["inst"] / tot_i, 100.0 * a["samp"] / max(tot_s, 1),
          a["thr"] / max(a["inst"], 1), " ".join("%8d" % a[k] for k in cols)))
  print("total inst", tot_i, "samples", tot_s)

if __name__ == "__main__":
  main(*sys.argv[1:4])
