"""Print the headline numbers of one bench.py JSON line read from stdin (helper for A/B runs on the GPU box)."""
import json
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else ""
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
bd = {k: v["us"] for k, v in (d.get("breakdown") or {}).items()}
print(tag, "img/s", round(d["value"]), "us/step", round(d["ms_per_step"] * 1e3, 2), "e2e", round(d["e2e"]["value"]), "launches", d.get("launches_per_step"), bd)
print("   clocks:", d.get("clocks"), "exchange:", (d.get("config") or {}).get("body_exchange"))
