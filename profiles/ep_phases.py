"""Critical-path breakdown of the persistent episode kernel: clock64 stamps of CTA 0 per phase."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = 60
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("phase_clocks", 1)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    pol.env_step(steps)
    pol.env_step(steps)
    c = pol.phase_clocks(steps)[5:]
names = ["W1 issue+fwd s,s' (to Z-safe barrier)", "target/dz1/dW2 (to bwd start)", "dW1 contraction + partial write",
         "window shift", "grid.sync #1", "phase B reduce+AdaGrad", "grid.sync #2"]
d = np.diff(c, axis=1).astype(np.float64)
tot = (c[1:, 0] - c[:-1, 0]).mean()
for k, nm in enumerate(names):
    print(f"{nm:45s} {d[:, k].mean():9.0f} cyc  {d[:, k].mean() / tot * 100:5.1f}%")
print(f"{'step total':45s} {tot:9.0f} cyc  (~{tot / 1.965e3:.1f} us at 1965 MHz)")
