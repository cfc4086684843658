"""Critical-path breakdown of episode_tc_kernel (clock64 stamps of CTA 0)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n, steps = (int(sys.argv[1]) if len(sys.argv) > 1 else 4096), 60
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("ep_tc", 1)
    pol.set_option("phase_clocks", 1)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    pol.env_step(steps)
    pol.env_step(steps)
    c = pol.phase_clocks(steps)[5:]
names = ["S0 operand build (A fwd, tables)", "S1 forward MMA (TMA ring + tcgen05)", "S2 epilogue 1 (h(s), z(s') out of TMEM)",
         "S3 select, book-keeping, Q(s')", "S4 TD, dZ1 hi/lo, h*dz2, X^T operand", "S5 backward MMA + dW2 + extra rows",
         "S6 epilogue 2 (dW1 partial out of TMEM)"]
d = np.diff(c[:, :11], axis=1).astype(np.float64)
tot = (c[1:, 0] - c[:-1, 0]).mean()
for k, nm in enumerate(names):
    print(f"{nm:48s} {d[:, k].mean():9.0f} cyc  {d[:, k].mean() / tot * 100:5.1f}%")
for k, nm in ((7, "S7 window shift"), (8, "grid.sync 1 (wait for the slowest CTA)"), (9, "phase B: sum + AdaGrad + W1^T publish")):
    print(f"{nm:48s} {d[:, k].mean():9.0f} cyc  {d[:, k].mean() / tot * 100:5.1f}%")
rest = (c[1:, 0] - c[:-1, 10]).mean()
print(f"{'grid.sync 2':48s} {rest:9.0f} cyc  {rest / tot * 100:5.1f}%")
print(f"{'step total':48s} {tot:9.0f} cyc  (~{tot / 1.965e3:.1f} us)")

# finer stamps (relative to the phase starts): S4 internals of warp 0, S6 of warp 11 and warp 0
f = lambda a, b: (c[:, a] - c[:, b]).mean()
print(f"phase B: all groups summed after {f(14, 9):.0f}; AdaGrad + publish {f(10, 15):.0f}")
if n > 148 * 32:
    print(f"last wave of CTA 0: carry load + window reload {f(13, 11):.0f}; S0 {f(1, 13):.0f}; S1..S6 {f(7, 1):.0f}")
