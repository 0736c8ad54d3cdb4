"""How accurate is the emulated compute? requested vs device-timed, single and chained spins."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlnetbench_b200 import capi

capi.init(0, 1)
a, b = capi.spin_calibrate()
print(json.dumps({"overhead_us": a, "slope": b}))
for ctas in (1, 148):
    capi.check(capi.lib().dlnb_set(b"spin_ctas", ctas))
    capi.spin_calibrate()
    for us, k in [(9, 100), (38, 100), (100, 10), (1000, 10), (2336, 10), (11682, 1), (35000, 1), (234562, 1)]:
        res = []
        for rep in range(3):
            capi.iter_begin()
            for _ in range(k):
                capi.compute_spin(us)
            ms = capi.iter_end()
            res.append(ms * 1000 / k)
        print(json.dumps({"spin_ctas": ctas, "req_us": us, "chain": k, "got_us_per_spin": [round(x, 3) for x in res]}))
capi.finalize()
