#!/usr/bin/env python
"""Reads bench.py's JSON line on stdin and prints the variant, MLUPS, ms/step and the per-kernel ms per launch."""
import json
import os
import sys

d = json.loads(sys.stdin.readlines()[-1])
k = d["roofline"]["kernels"]
print(os.environ.get("LBM_PF_BLOCKS", ""), d["config"]["kernel_variant"], round(d["value"]), round(d["ms_per_step"], 3),
      {a: round(b["ms"] / b["launches"], 3) for a, b in k.items()})
