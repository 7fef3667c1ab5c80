import json, sys
src = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
line = [l for l in src.read().splitlines() if l.startswith("{")][-1]
d = json.loads(line)
r = d.get("roofline") or {}
print(sys.argv[1].split('/')[-1] if len(sys.argv) > 1 else "-", "val %.0f e2e %.0f roof %.0f %s (%.2f)" % (
    d["value"], d["e2e"]["value"], r.get("achieved", 0), r.get("unit", ""), r.get("frac", 0)), d.get("stages_ms"))
