import json, sys
for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l); print(d["clocks"], "value %.0f" % d["value"], "n_gpus", d["n_gpus"])
