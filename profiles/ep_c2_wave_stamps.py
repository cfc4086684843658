import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np
import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O
n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
steps = 12
prices = O.gen_prices(42, 0, n, 300)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("phase_clocks", 1)
    pol.set_option("fanout", 0)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    assert pol.get_option("ep_c2_active") == 1
    pol.env_step(steps)
    c = pol.phase_clocks(steps)[3:]
order = [0, 11, 13, 1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 12, 10]
print(n, "episodes; step total", (c[1:, 0] - c[:-1, 0]).mean())
for a, b in zip(order[:-1], order[1:]):
    print(f"  slot {a:2d} -> {b:2d}: {(c[:, b] - c[:, a]).mean():9.0f}")
