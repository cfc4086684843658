"""configs[2]/configs[4] per-GPU shape: 131 072 episodes on one GPU (512 tickers x 256 episodes each, or
131 072 distinct series): the fan-out path (default where the episodes are grouped by series), the tensor-core
persistent kernel in waves and the three-kernels-per-step path.  Usage: python profiles/fanout_workload.py [tickers] [steps]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

tickers = int(sys.argv[1]) if len(sys.argv) > 1 else 512
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
n = 131072
prices = O.gen_prices(42, 0, tickers, 1000)
soe = (np.arange(n) // (n // tickers)).astype(np.int32)
for name, opts in (("default", {}), ("pair kernel in waves", {"fanout": 0}),
                   ("one-tile-per-CTA tensor-core kernel in waves", {"fanout": 0, "ep_c2": 0}),
                   ("three kernels per step", {"fanout": 0, "fused": 0})):
    with q.QDecisionPolicy(q.PolicyConfig()) as pol:
        for k, v in opts.items():
            pol.set_option(k, v)
        pol.env_create(prices, soe, 1000.0, 0)
        act = {k: int(pol.get_option(k)) for k in ("fanout_active", "fused_active", "ep_tc_active", "ep_c2_active")}
        pol.env_step(10)
        t0 = time.perf_counter()
        pol.env_step(steps)
        dt = time.perf_counter() - t0
        kms = pol.env_step_timed(steps) / steps * 1e3
        print(f"{name} {act}: tickers {tickers} episodes {n}: "
              f"{dt / steps * 1e6:.1f} us/env-step -> {n * steps / dt:.4g} env-steps/s; kernels us {kms.tolist()}")
