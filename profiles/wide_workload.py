"""Short workload for ncu: configs[3] (window 513, hidden 2048, 65 536 episodes), a few env-steps on the
GEMM-organised path.  Usage: python profiles/wide_workload.py [tc=1|0] [steps] [passes]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

tc = int(sys.argv[1]) if len(sys.argv) > 1 else 1
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 3
n = 65536
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig(window=513, hidden=2048)) as pol:
    pol.set_option("tc", tc)
    pol.set_option("tc_passes", passes)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    pol.env_step(steps)
    ms = pol.env_step_timed(steps)[0] / steps
    print("ok", pol.steps_done, pol.launch_count, "tc_active", pol.get_option("tc_active"), "ms/env-step", ms)
