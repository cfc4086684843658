"""Config 2 (4096 episodes): persistent kernel, FFMA vs tcgen05 contractions.  Usage: python profiles/ep_tc_timing.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

n = 4096
prices = O.gen_prices(42, 0, n, 1000)
for tc in (0, 1):
    with q.QDecisionPolicy(q.PolicyConfig()) as pol:
        pol.set_option("ep_tc", tc)
        pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
        assert pol.get_option("ep_tc_active") == tc
        pol.env_step(100)
        ms = pol.env_step_timed(600)[0]
        print(f"ep_tc={tc}: {ms / 600 * 1e3:.2f} us per env-step of {n} episodes -> {n * 600 / (ms * 1e-3):.4g} env-steps/s")
