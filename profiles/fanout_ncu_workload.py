import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np
import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O
n, tickers = 131072, 512
prices = O.gen_prices(42, 0, tickers, 1000)
soe = (np.arange(n) // (n // tickers)).astype(np.int32)
with q.QDecisionPolicy(q.PolicyConfig()) as pol:
    pol.env_create(prices, soe, 1000.0, 0)
    assert pol.get_option("fanout_active") == 1
    pol.env_step(6)
    print("ok")
