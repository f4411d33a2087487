"""Prints the headline fields of the last JSON line on stdin (scratch helper of the tools/gpu_*.sh scripts)."""
import json
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else ""
lines = [l for l in sys.stdin.read().strip().split("\n") if l.startswith("{")]
d = json.loads(lines[-1])
print(tag, "value", round(d["value"]), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"]),
      "ratio", round(d["e2e"]["value"] / d["value"], 4), d.get("clocks"))
