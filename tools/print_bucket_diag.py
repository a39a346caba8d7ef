import json
import sys

for line in open(sys.argv[1]):
    if line.startswith("{"):
        d = json.loads(line)
        print("world", d["world"], "ms_per_step", round(d["ms_per_step"], 3))
        for mode, rows in (d.get("bucket_kernels_alone") or {}).items():
            print(" ", mode, [(r["layer"], r["us"]) for r in rows])
