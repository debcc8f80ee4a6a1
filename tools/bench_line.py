"""Print a compact summary of one bench.py JSON line read from stdin (label = argv[1])."""
import json
import sys

d = json.loads(sys.stdin.read())
big = " ".join(f"{k.split('(')[-1].rstrip(')')[:9]}={v['ms'] * 1000:.1f}" for k, v in d["kernels"].items() if v["ms"] > 0.02)
print(sys.argv[1] if len(sys.argv) > 1 else "", round(d["ms_per_step"], 4), "ms", round(d["value"] / 1e6, 3), "M/s e2e",
      round(d["e2e"]["value"] / 1e6, 3), "|", big)
