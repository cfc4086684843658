"""Per-CTA globaltimer stamps of the pair kernel ("phase_clocks" = 2): who arrives when at the two grid-wide waits."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n, steps, G = (int(sys.argv[1]) if len(sys.argv) > 1 else 4096), 40, 148
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("phase_clocks", 2)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    assert pol.get_option("ep_c2_active") == 1
    pol.env_step(steps)
    out = np.zeros((steps, G, 8), np.int64)
    q.policy.check(pol._lib.qp_debug_phase_clocks(pol._h, steps * G * 8 // 16, out.ctypes.data_as(C.c_void_p)))
t = out[5:].astype(np.float64)
tiles = 2 * ((n + 63) // 64)
ns = np.diff(t[:, 0, 0]).mean()
print(f"{n} episodes: step {ns:.0f} ns; {tiles} CTAs hold tiles")
names = ["step top", "weights seen, first W tiles requested", "forward done", "dZ1 written", "partial stored (arrives at grid barrier)",
         "grid barrier passed", "partials summed", "phase B done"]
t0 = t[:, :, 0].min(axis=1, keepdims=True)
for i, nm in enumerate(names):
    sel = slice(0, tiles) if i in (1, 2, 3) else slice(0, G)
    d = (t[:, sel, i] - t0).mean(axis=0)
    print(f"  {nm:44s} mean {d.mean():7.0f} ns  first {d.min():7.0f} (#{d.argmin()})  last {d.max():7.0f} (#{d.argmax()})  p10 {np.percentile(d, 10):7.0f}  p90 {np.percentile(d, 90):7.0f}")
dur = lambda a, b, sel: (t[:, sel, b] - t[:, sel, a]).mean(axis=0)
for a, b, nm in ((1, 2, "forward"), (2, 3, "element-wise"), (3, 4, "backward + partial out"), (5, 6, "sum"), (6, 7, "AdaGrad + publish")):
    sel = slice(0, tiles) if a in (1, 2, 3) else slice(0, G)
    d = dur(a, b, sel)
    print(f"  duration {nm:26s} mean {d.mean():7.0f}  min {d.min():7.0f} (#{d.argmin()})  max {d.max():7.0f} (#{d.argmax()})")
