"""Short workload for ncu: config 2 with the pair kernel (episode_c2_kernel), two 30-step launches."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n = 4096
prices = O.gen_prices(42, 0, n, 1000)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
    assert pol.get_option("ep_c2_active") == 1
    pol.env_step(30)
    pol.env_step(30)
    print("ok", pol.steps_done, pol.get_option("ep_c2_active"))
