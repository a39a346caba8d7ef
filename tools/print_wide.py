import json
import sys

for line in sys.stdin:
    if line.startswith("{"):
        d = json.loads(line)
        print(d["hidden"], round(d["ms_per_step"], 4), {k: round(v * 1e3, 1) for k, v in d["kernels_ms"].items()})
