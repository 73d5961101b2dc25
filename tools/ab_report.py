#!/usr/bin/env python3
"""Tabulate gpurun_out/<tag>_v<variant>_s<scene>[_fast].json written by tools/ab_variants.sh."""
import glob, json, os, re, sys
tag = sys.argv[1] if len(sys.argv) > 1 else "ab"
rows = {}
for f in glob.glob(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", f"{tag}_v*_s*.json")):
    m = re.search(rf"{tag}_v(-?\d+)_s(\d+)(_fast)?\.json", f)
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        ms = min(d["plain_dispatch_ms"])
    except Exception:
        ms = float("nan")
    rows.setdefault((int(m.group(2)), bool(m.group(3))), {})[int(m.group(1))] = ms
for key in sorted(rows):
    print(f"scene {key[0]}{' fast' if key[1] else '     '}:", "  ".join(f"v{v}: {ms:.3f}" for v, ms in sorted(rows[key].items())))
