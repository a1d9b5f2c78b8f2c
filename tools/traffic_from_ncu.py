#!/usr/bin/env python
"""Turns an ncu launch list (--csv, metrics dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum) of
`bench.py --steps N` at 512^3 into profiles/traffic.json: DRAM bytes per launch for every kernel of the step, keyed by
the name the library's profiler reports (lbm_profile_report), with the full template names kept so that bench.py can
tell which kernels the capture is of.

    python tools/traffic_from_ncu.py gpurun_out/.../traffic.csv "command line of the capture" > profiles/traffic.json
"""
import collections
import csv
import json
import re
import sys


def main():
    path, how = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    lines = open(path).read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    launches = collections.OrderedDict()
    for r in csv.DictReader(lines[start:]):
        d = launches.setdefault(r["ID"], {"name": r["Kernel Name"], "grid": r.get("Grid Size", "")})
        d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    per = collections.defaultdict(list)
    for d in launches.values():
        m = re.search(r"(k_[a-z_0-9]+)(<[^>]*>)?", d["name"])
        base, targs = m.group(1), m.group(2) or ""
        if base == "k_step_one" and targs.endswith(", 0>"):
            base = "k_collide_core"      # k_step_one<.., PHI = false>: the name the library's profiler gives the collide-only launches
        # output-step flavours (<FXY, WRITE=1, ...>) also write the four macroscopic fields: one step in a hundred
        if base.startswith("k_collide") and re.match(r"<\d+, 1", targs):
            continue
        per[base].append((d.get("dram__bytes_read.sum", 0.0), d.get("dram__bytes_write.sum", 0.0), d.get("gpu__time_duration.sum", 0.0),
                          base + targs, d["grid"]))
    out = {"_source": how, "_unit": "bytes per launch, mean over the captured launches; ncu times are cold-cache and serialised",
           "grid_nodes": 512 ** 3, "kernels": {}}
    for base, rows in per.items():
        n = len(rows)
        out["kernels"][base] = {"template": sorted({r[3] for r in rows}), "launches_captured": n,
                                "dram_bytes_read_per_launch": sum(r[0] for r in rows) / n,
                                "dram_bytes_write_per_launch": sum(r[1] for r in rows) / n,
                                "dram_bytes_per_launch": sum(r[0] + r[1] for r in rows) / n,
                                "ncu_ns_per_launch": sum(r[2] for r in rows) / n, "grid": sorted({r[4] for r in rows})}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
