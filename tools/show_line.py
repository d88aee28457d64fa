"""Print the key fields of a bench.py JSON line (file argument)."""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("ms/step", round(d["ms_per_step"], 3), {k: round(v, 3) for k, v in d["phase_ms_per_step"].items()}, "value", f'{d["value"]:.4g}',
      "launches/step", d["gpu_launches"] / d["steps"], "roofline", round(d["roofline"]["frac"], 3), "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
if d.get("e2e"):
    print("e2e", f'{d["e2e"]["value"]:.4g}', round(d["e2e"]["ms_per_step"], 3), "ms/step over", d["e2e"]["steps"], "steps")
if d.get("strong_scaling"):
    s = d["strong_scaling"]
    print("strong", round(s["ms_per_step"], 3), {k: round(v, 3) for k, v in s["phase_ms_per_step"].items()}, f'{s["value"]:.4g}')
for k, v in d.get("workloads", {}).items():
    print(" ", k, round(v["ms_per_step"], 2), {a: round(b, 2) for a, b in v["phase_ms_per_step"].items()}, v.get("gpu_launches_per_step"),
          round(v.get("roofline", {}).get("frac") or 0, 3))
