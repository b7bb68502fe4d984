#!/usr/bin/env python
"""print an exchange_sweep.py result (the JSON may be preceded by library banner lines)"""
import json
import sys

for path in sys.argv[1:]:
    try:
        text = open(path).read()
        d = json.loads(text[text.index("{"):])
    except Exception as e:       # noqa: BLE001
        print(path, "no result", e)
        continue
    print("%s: %d GPUs, symmetric memory %s: fused %.0f us total, NCCL all-reduce + update %.0f us" % (
        path, d["n_gpus"], d["symmetric_memory"], d["total_fused_us"], d["total_nccl_us"]))
    for r in d["rows"]:
        print("  %-6s %8.2f MB  fused %8.1f us  %6.0f GB/s per direction (%.2f of 770)   nccl %8.1f us (%.0f GB/s)" % (
            r["layer"], r["MB"], r["fused_us"], r["fused_GBps_per_dir"], r["frac_of_770GBps"],
            r["nccl_allreduce_update_us"], r["nccl_GBps_per_dir"]))
