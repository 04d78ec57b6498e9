import json, sys
for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l)
        print({k: d[k] for k in d if k in ("ms_per_cycle", "ms_per_step", "value", "single_domain_graph")})
