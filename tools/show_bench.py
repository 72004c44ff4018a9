#!/usr/bin/env python3
"""One line per (sub-)record of a bench.py JSON line: value, roofline fraction, kernel time, e2e."""
import json
import sys

line = json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][-1])


def show(r, name):
    e = r.get("e2e") or {}
    rf = r.get("roofline") or {}
    print("%-12s value %9.1f M  frac %.3f  kernel %.4f ms  e2e %9.1f M" % (name, r["value"] / 1e6, rf.get("frac", 0), rf.get("kernel_ms", 0), e.get("value", 0) / 1e6))


show(line, line["config"]["workload"])
for k, v in line.get("sub_records", {}).items():
    show(v, k)
print("clocks", line.get("clocks"), "| cpu_baseline %.2f M" % ((line.get("cpu_baseline") or {}).get("value", 0) / 1e6), "| n_gpus", line.get("n_gpus"),
      "| weights", line.get("weights_sha256_16"))
