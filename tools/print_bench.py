import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
us = lambda v: round(v * 1e3, 2)
print("n", d["n_gpus"], "value us", us(d["ms_per_step"]), "hot", us(d["hot_l2"]["ms_per_step"]), "e2e", us(d["e2e"]["ms_per_step"]),
      "e2e_gloo", d["e2e_gloo_loss"] and us(d["e2e_gloo_loss"]["ms_per_step"]), "weak", d["weak"] and round(d["weak"]["value"] / 1e6, 1))
print("parity", {k: (v.get("ok"), v.get("loss_rel_max"), v.get("param_abs_max")) for k, v in d["parity"].items() if isinstance(v, dict) and "ok" in v},
      "all", d["parity"]["ok"])
w = d["wide"]
print("wide ms", round(w["ms_per_step"], 4), "buckets", w["buckets"], "weak", w.get("weak"), "speedup", w["speedup_vs_incumbent"])
print("pipeline", d["pipeline"] and (d["pipeline"]["ms_per_step_device"], d["pipeline"]["ms_per_step_with_loss_read"], d["pipeline"]["speedup_vs_incumbent"]))
print("value_per_step", d.get("value_per_step"))
print("clocks", d["clocks"], "launches", d["gpu_launches"])
