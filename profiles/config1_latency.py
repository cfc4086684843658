"""configs[0]-shaped run on the GPU: 1 episode, one 1000-step series, 799 sequential env-steps through
the persistent kernel (B = 1 is exactly the reference's per-message update)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

prices = O.gen_prices(42, 0, 1, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.env_create(prices, [0], 1000.0, 0)
    pol.env_step(799)
    for n in (1, 10):
        best = 1e9
        for _ in range(5):
            pol.env_reset()
            t0 = time.perf_counter()
            pol.env_step(799)
            best = min(best, time.perf_counter() - t0)
    print(f"1 episode x 799 env-steps: {best * 1e3:.2f} ms -> {best / 799 * 1e6:.1f} us per env-step, {799 / best:.4g} env-steps/s")
o = O.Oracle(O.Config())
o.env_create(prices, [0], 1000.0, 0)
t0 = time.perf_counter()
o.env_step(799, mode=1)
dt = time.perf_counter() - t0
print(f"CPU oracle (1 core): {dt * 1e3:.1f} ms -> {799 / dt:.4g} env-steps/s")
