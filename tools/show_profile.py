#!/usr/bin/env python3
"""Print the lines of tools/profile_day.py output files compactly: start day, us per day, per-kernel us per day."""
import json
import sys

for f in sys.argv[1:]:
    print("#", f)
    for l in open(f):
        try:
            d = json.loads(l)
        except Exception:
            if l.strip():
                print(l[:200].rstrip())
            continue
        k = d.get("kernels_us_per_day")
        print(d["start_day"], d["days"], d["device_us_per_day"], " ".join(f"{a[2:]}={b:.0f}" for a, b in k.items()) if k else "", "| act", d["work_per_day"]["active"], "trials", d["work_per_day"]["trials"])
