#!/usr/bin/env python
"""Print the handful of numbers of a bench.py JSON line that a tuning pass looks at."""
import json, sys
d = json.load(open(sys.argv[1]))
e, r = d["e2e"], d["roofline"]
print(f"value {d['value']/1e6:.3f} M ({d['ms_per_step']*1e3:.1f} us)  e2e {e['value']/1e6:.3f} M ({e['ms_per_step']*1e3:.1f} us, d2h {e['d2h_bytes_per_step']} B)")
if d.get("worst_case"):
    w = d["worst_case"]
    print(f"worst {w['value']/1e6:.3f} M ({w['ms_per_step']*1e3:.1f} us, ratio {w['ratio_to_value']:.3f})  e2e {w['e2e']['value']/1e6:.3f} M ({w['e2e']['ms_per_step']*1e3:.1f} us)")
if d.get("sustained"):
    s = d["sustained"]
    print(f"sustained {s['value']/1e6:.3f} M ({s['ms_per_step']*1e3:.1f} us over {s['seconds']:.1f} s, ratio {s['ratio_to_value']:.3f}) clocks {s['clocks']['sm_mhz']} {s['clocks']['reasons']} n={s['clocks']['samples']}")
print(f"bank {r['kernel_ms']*1e3:.1f} us frac {r['frac']:.3f} chain_frac {r.get('chain_frac', 0):.3f} serialised block {r['block_ms_serialised']*1e3:.1f} us; host us {e['host_us_per_step']}")
for k in ("channel_sharded", "config4"):
    if d.get(k):
        print(k, json.dumps(d[k])[:600])
