"""Prints the headline numbers and per-entry-point shares of a bench.py JSON line.  python tools/bench_summary.py file.json"""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms/step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), "eval", round(d.get("eval", {}).get("value", 0)),
      "roofline", d.get("roofline", {}).get("frac"))
for k, v in d.get("kernel_shares", {}).items():
    print(f"  {k:34s} {v['ms']:.4f} ms  x{v['launches']}")
