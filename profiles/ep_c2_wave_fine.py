"""Fine stamps of the last tile of a wave-mode step of the pair kernel (CTA 0): python profiles/ep_c2_wave_fine.py [episodes]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
steps = 12
prices = O.gen_prices(42, 0, n, 300)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("phase_clocks", 3)
    pol.set_option("fanout", 0)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    assert pol.get_option("ep_c2_active") == 1
    pol.env_step(steps)
    out = np.zeros((steps, 64), np.int64)
    q.policy.check(pol._lib.qp_debug_phase_clocks(pol._h, steps, out.ctypes.data_as(C.c_void_p)))
c = out[3:]
t48 = c[:, 48]
rel = lambda a: round((c[:, a] - t48).mean())
print("relative to the start of the tile's phase A (slot 48):")
print("  W tile kk requested :", [rel(k) for k in range(7)])
print("  MMA: A tile kk ready:", [rel(16 + k) for k in range(7)])
print("  MMA: W tile kk ready:", [rel(8 + k) for k in range(7)])
print("  last commit", rel(40), " forward complete", rel(41))
