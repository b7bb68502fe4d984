#!/usr/bin/env python
"""print the interesting keys of a bench.py JSON line (last line of the given file or stdin)"""
import json
import sys

src = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
lines = [l.strip() for l in src if l.strip()]
for l in lines[:-1]:
    if "Warn" in l or "warn" in l or "rror" in l:
        print(l[:300])
d = json.loads(lines[-1])
print(d["n_gpus"], "gpu", round(d["value"]), "img/s", round(d["ms_per_step"], 3), "ms/step; host", round(d.get("host_enqueue_ms_per_step", 0), 3),
      "ms; launches", d.get("gpu_launches"), "; e2e", round(d["e2e"]["value"]), "img/s", round(d["e2e"]["ms_per_step"], 3), "ms")
