"""clock64 stamps of CTA 0 of the persistent tensor-core kernels: python profiles/ep_stamps.py <episodes> <c2|tc> [steps]
Slots: 0 step start, 1 split barrier passed (weights final, first W tiles requested), 2 forward MMA complete, 3 z in shared
memory, 4 select done, 5 dZ1 written, 6 backward MMA complete, 7 partial handed to TMA, 8 stores complete, 9 grid barrier
passed, 14 partial segments summed, 15 (same, last pass), 12 NVLink exchange done, 10 AdaGrad + publish done."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
path = sys.argv[2] if len(sys.argv) > 2 else "c2"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 60
dbg = int(sys.argv[4]) if len(sys.argv) > 4 else 0
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("ep_c2", 1 if path == "c2" else 0)
    pol.set_option("phase_clocks", 1)
    if dbg:
        pol.set_option("c2_dbg", dbg)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    assert pol.get_option("ep_c2_active") == (1 if path == "c2" else 0) and pol.get_option("ep_tc_active") == 1
    pol.env_step(steps)
    pol.env_step(steps)
    c = pol.phase_clocks(steps)[5:]
    ms = pol.env_step_timed(200)[0] / 200
order = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 12, 10]
tot = (c[1:, 0] - c[:-1, 0]).mean()
print(f"{path} dbg={dbg} {n} episodes: {ms * 1e3:.2f} us per env-step (CUDA events), {tot:.0f} cycles per step on CTA 0")
for a, b in zip(order[:-1], order[1:]):
    if (c[:, a] == 0).all() or (c[:, b] == 0).all():
        continue
    print(f"   slot {a:2d} -> {b:2d}: {(c[:, b] - c[:, a]).mean():8.0f}")
print(f"   slot 10 -> next 0: {(c[1:, 0] - c[:-1, 10]).mean():8.0f}")
