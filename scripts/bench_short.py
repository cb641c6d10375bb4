"""Condensed view of bench.py's JSON line (reads stdin)."""
import json, sys
for line in sys.stdin:
    try:
        d = json.loads(line)
    except Exception:
        print(line[:300].rstrip()); continue
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"])
    r = d["roofline"]
    print({k: r[k] for k in ("achieved", "peak", "frac", "force_ms_per_launch", "ordered_pairs_in_cutoff", "candidate_pairs_evaluated")})
    m = d["roofline_mech"]; print({k: m.get(k) for k in ("achieved", "frac", "ms_per_step")})
    if "cpu_baseline" in d: print(d["cpu_baseline"])
