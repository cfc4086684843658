"""Stamps inside the forward of the pair kernel (CTA 0): python profiles/ep_fine.py <episodes> [dbg]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dbg = int(sys.argv[2]) if len(sys.argv) > 2 else 0
steps = 40
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("phase_clocks", 3)
    pol.set_option("c2_dbg", dbg)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    assert pol.get_option("ep_c2_active") == 1
    pol.env_step(steps)
    out = np.zeros((steps, 64), np.int64)
    q.policy.check(pol._lib.qp_debug_phase_clocks(pol._h, steps, out.ctypes.data_as(C.c_void_p)))
c = out[5:]
t0 = c[:, 0:1]        # producer: first W tile requested
rel = lambda a: (c[:, a] - t0[:, 0]).mean()
print(f"episodes {n} dbg {dbg}: cycles after the first W request (mean over {len(c)} steps)")
print("  W tile kk requested   :", [round(rel(k)) for k in range(7)])
print("  A tile kk free (bw 0) :", [round(rel(24 + k)) for k in range(7)])
print("  A tile kk built (bw 0):", [round(rel(32 + k)) for k in range(7)])
print("  MMA: A tile kk ready  :", [round(rel(16 + k)) for k in range(7)])
print("  MMA: W tile kk ready  :", [round(rel(8 + k)) for k in range(7)])
print("  last commit issued    :", round(rel(40)), "  forward complete seen by thread 0:", round(rel(41)))

d = lambda a, b: round((c[:, b] - c[:, a]).mean())
print("  element-wise (thread 0): S2", d(41, 39), " S3 loop", d(39, 42), " barrier", d(42, 43), " reduce + peer store", d(43, 44),
      " cluster barrier", d(44, 45), " select + dZ1 loop", d(45, 46), " rest of S4", d(46, 47), " barrier", d(47, 51))
print("  S6 (thread 0): combine", d(52, 53), " boxes", d(53, 54), " next price", d(54, 55), " barrier", d(55, 58), "; warp 11 boxes", d(56, 57))
