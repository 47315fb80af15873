"""Aggregates an ncu source page (cuda view) by file:line; prints the top lines.
usage: ncu_by_line.py report.ncu-rep [N]"""
import csv, subprocess, sys

def main(rep, top=40):
  out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                       capture_output=True, text=True).stdout
  rows = list(csv.reader(out.splitlines()))
  fname, hdr, agg = "?", None, []
  for r in rows:
    if len(r) == 2 and r[0] == "File Path":
      fname = r[1].split("/")[-1]; continue
    if len(r) == 2:
      continue
    if r and r[0] == "Line No":
      hdr = r; continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
      continue
    try:
      agg.append((fname, int(r[0]), r[1].strip()[:90], int(r[hdr.index("# Samples")]),
                  int(r[hdr.index("Instructions Executed")]),
                  int(r[hdr.index("stall_no_inst")]), int(r[hdr.index("stall_long_sb")])))
    except ValueError:
      continue
  ts = sum(a[3] for a in agg); ti = sum(a[4] for a in agg)
  print("total samples %d, instructions %d" % (ts, ti))
  for a in sorted(agg, key=lambda a: -a[3])[:top]:
    print("%-18s %4d  samp %5.2f%%  inst %5.2f%%  noinst %6d longsb %6d | %s" %
          (a[0], a[1], 100.0 * a[3] / ts, 100.0 * a[4] / ti, a[5], a[6], a[2]))

if __name__ == "__main__":
  main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
