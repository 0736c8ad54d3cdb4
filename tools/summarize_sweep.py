import json, sys
for f in sys.argv[1:]:
    rows = json.load(open(f))
    print(f)
    for r in rows:
        print(f"  {r['op']:14s} w{r['world']} {r['bytes']:>11d} B ctas={r['ctas']:<4d} batch {r['us_b2b']:9.1f} us  slot {r['us_slot_med']:9.1f} us  busBW {r['busbw_GBps']:8.1f} GB/s")
