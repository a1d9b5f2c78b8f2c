#!/usr/bin/env python
"""Summarises `ncu -i X.ncu-rep --page source --csv --print-source sass`: samples and executed instructions per
instruction, the most-sampled instructions, and stall-reason totals.  Usage: ncu_sass_profile.py file.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
col = {n: i for i, n in enumerate(hdr)}
data = [r for r in rows[hdr_i + 1:] if len(r) >= len(hdr) - 2]


def f(r, name):
    try:
        return float(r[col[name]])
    except Exception:
        return 0.0


tot_s = sum(f(r, "# Samples") for r in data)
tot_i = sum(f(r, "Instructions Executed") for r in data)
print("instructions (static): %d   warp instructions executed: %.4g   samples: %d" % (len(data), tot_i, tot_s))
stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
print("stall totals:", ", ".join("%s %.1f%%" % (n[6:], 100 * sum(f(r, n) for r in data) / max(tot_s, 1)) for n in
                                 sorted(stalls, key=lambda n: -sum(f(r, n) for r in data))[:8]))
print("\ncumulative profile in 20 chunks of the instruction stream: [first..last] executed%  samples%")
n = len(data)
for c in range(20):
    part = data[c * n // 20:(c + 1) * n // 20]
    print("  %4d..%4d  exec %5.1f%%  samples %5.1f%%   %s" % (c * n // 20, (c + 1) * n // 20 - 1, 100 * sum(f(r, "Instructions Executed") for r in part) / tot_i,
                                                             100 * sum(f(r, "# Samples") for r in part) / max(tot_s, 1), part[0][col["Source"]].strip()[:40]))
print("\ntop instructions by samples:")
for i, r in sorted(enumerate(data), key=lambda t: -f(t[1], "# Samples"))[:top]:
    why = max(stalls, key=lambda s: f(r, s))
    print("  #%4d %5.2f%%  exec %.3g  %-14s %s" % (i, 100 * f(r, "# Samples") / max(tot_s, 1), f(r, "Instructions Executed"), why[6:], r[col["Source"]].strip()[:70]))
