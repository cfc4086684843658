"""Prices that are not tf32-exact (quarter ticks x 1.0371, as cent quotes are not): time per env-step of the pair kernel
in its split form (32 real + 32 virtual episodes per pair) against the FFMA persistent kernel and the three-kernel path.
    python profiles/ep_c2_split.py [episodes ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

sizes = [int(a) for a in sys.argv[1:]] or [10, 512, 2368, 4096]
for n in sizes:
    prices = O.gen_prices(42, 0, n, 1000) * np.float32(1.0371)
    row = []
    for name, opts in (("default", {}), ("ffma", {"ep_tc": 0}), ("three", {"fused": 0})):
        with q.QDecisionPolicy(q.PolicyConfig()) as pol:
            for k, v in opts.items():
                pol.set_option(k, v)
            pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
            act = {k: int(pol.get_option(k)) for k in ("fused_active", "ep_tc_active", "ep_c2_active", "ep_c2_split_active")}
            pol.env_step(40)
            steps = 200 if act["fused_active"] else 50
            us = float(np.sum(pol.env_step_timed(steps))) / steps * 1e3      # (all kernels of a step: the three-kernel path fills three slots)
            row.append(f"{name}: {us:7.2f} us/step ({n / us * 1e6:.3g} env-steps/s) {act}")
    print(f"{n} episodes, inexact prices")
    for r in row:
        print("   ", r)
