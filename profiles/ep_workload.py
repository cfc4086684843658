"""Short workload for ncu: config 2 (4096 episodes), two launches of the persistent episode kernel
(30 env-steps each).  Usage: python profiles/ep_workload.py [fused=1|0] [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

fused = int(sys.argv[1]) if len(sys.argv) > 1 else 1
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
n = 4096
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.set_option("fused", fused)
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    pol.env_step(steps)
    pol.env_step(steps)
    print("ok", pol.steps_done, pol.launch_count)
