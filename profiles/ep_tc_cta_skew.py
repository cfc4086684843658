"""Which CTAs of episode_tc_kernel are on the critical path: globaltimer stamps of EVERY CTA ("phase_clocks" = 2)
at the step start, after the forward MMA, before the grid barrier and after phase B."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n, steps, G = 4096, 40, 148
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("phase_clocks", 2)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    pol.env_step(steps)
    pol.env_step(steps)
    raw = pol.phase_clocks(steps * G * 4 // 16)
t = raw.reshape(steps, G, 4)[5:].astype(np.float64)
tiles = n // 32
step_ns = np.diff(t[:, 0, 0]).mean()
print(f"step {step_ns:.0f} ns")
t0 = t[:, :tiles, 0].min(axis=1, keepdims=True)           # earliest CTA start of the step
for i, name in ((0, "step start"), (1, "forward MMA done"), (2, "partial stored (arrives at the grid barrier)")):
    d = t[:, :tiles, i] - t0
    m = d.mean(axis=0)
    print(f"{name:46s} mean {m.mean():7.0f} ns   fastest CTA {m.min():7.0f} (#{m.argmin()})   slowest {m.max():7.0f} (#{m.argmax()})   p90 {np.percentile(m, 90):7.0f}")
d = (t[:, :, 3] - t0)
m = d.mean(axis=0)
print(f"{'phase B done (all CTAs)':46s} mean {m.mean():7.0f} ns   fastest {m.min():7.0f} (#{m.argmin()})   slowest {m.max():7.0f} (#{m.argmax()})")
arr = (t[:, :tiles, 2] - t0).mean(axis=0)
order = np.argsort(arr)
print("slowest ten CTAs to the barrier:", [(int(c), int(arr[c])) for c in order[-10:]])
print("fastest ten:", [(int(c), int(arr[c])) for c in order[:10]])
fwd = (t[:, :tiles, 1] - t[:, :tiles, 0]).mean(axis=0)
rest = (t[:, :tiles, 2] - t[:, :tiles, 1]).mean(axis=0)
print(f"start -> forward done: min {fwd.min():.0f} max {fwd.max():.0f}; forward done -> partial stored: min {rest.min():.0f} max {rest.max():.0f}")
print("corr(arrival, start->forward)", np.corrcoef(arr, fwd)[0, 1], " corr(arrival, kb_first = c % 7):", np.corrcoef(arr, np.arange(tiles) % 7)[0, 1])
