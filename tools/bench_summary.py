#!/usr/bin/env python
"""Print value / e2e and the per-kernel-group times of bench.py JSON lines: tools/bench_summary.py file.json [...]"""
import json
import sys

for f in sys.argv[1:]:
    j = json.load(open(f))
    print(f"{f}: {j['value']:.2f} {j['unit']}  {j['ms_per_step']:.1f} ms/step  e2e {j['e2e']['value']:.2f} ({j['e2e'].get('ms_per_step', 0):.1f} ms)  "
          f"launches {j['gpu_launches']}  groups {(j.get('result') or j['config']).get('ngroups')}  hash {(j.get('result') or {}).get('membership_hash')}  marginal {(j.get('result') or {}).get('n_marginal')}")
    r = j.get("roofline_all") or {}
    for k, v in sorted(r.items(), key=lambda kv: -kv[1].get("ms", 0))[:40]:
        extra = "  ".join(f"{a}={v[a]:.3g}" for a in ("achieved", "frac") if a in v)
        print(f"   {k:28s} {v['ms']:9.3f} ms  {extra}")
