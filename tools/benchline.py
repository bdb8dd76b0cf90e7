"""Reads bench.py's JSON line from stdin and prints the few numbers compared during kernel work."""
import json
import sys

d = json.loads(sys.stdin.read().strip().splitlines()[-1])
c3 = d.get("c3_frame0", {})
print("variant", d["config"].get("variant"), "| Mrays/s", round(d["value"]), "| ms/step", round(d["ms_per_step"], 3),
      "| kernel_ms", round(d.get("kernel_ms", 0), 3), "| e2e", round(d.get("e2e", {}).get("value", 0)), "| C3 frame0 ms", round(c3.get("ms_per_frame", 0), 3),
      "| launches", d.get("gpu_launches"), "| fp32 frac", round(d.get("roofline_fp32", {}).get("frac", 0), 3))
