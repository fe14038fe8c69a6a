"""Reads bench.py's JSON line from stdin and prints ms/step, value and the phase times."""
import json, sys
for l in sys.stdin:
    if l.startswith("{"):
        j = json.loads(l)
        d = j.get("detail", {})
        print(round(j["ms_per_step"], 1), "%.3e" % j["value"], "e2e %.3e" % j["e2e"]["value"],
              {k: round(v, 2) for k, v in d.get("phase_ms", {}).items()})
