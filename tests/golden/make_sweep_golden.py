#!/usr/bin/env python
"""Writes tests/golden/sweep_manifest.json from the reference's own stored sweep manifest
(/root/reference/outputs/05_grid_search/manifest.json): the 78 (experiment_id, cfg, steps, phase) rows the
reference actually generated.  Run in the build container (the reference is not on the GPU box)."""
import json
import os

SRC = "/root/reference/outputs/05_grid_search/manifest.json"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sweep_manifest.json")
m = json.load(open(SRC))
rows = [[e["experiment_id"], e["video_name"], e["cfg"], e["steps"], e["phase"]] for e in m["experiments"]]
json.dump({"source": SRC, "grid_params": m["grid_params"], "experiments": rows}, open(DST, "w"), indent=0)
print(len(rows), "runs,", sum(r[3] for r in rows), "denoise steps")
