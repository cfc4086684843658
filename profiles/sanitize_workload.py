"""Smoke-size workload for compute-sanitizer: every form of qp_env_step on a handful of episodes and env-steps.
    compute-sanitizer --tool memcheck python profiles/sanitize_workload.py [forms ...]
forms: c2 (pair kernel), c2s (pair kernel, prices not tf32-exact), tc (one tile per CTA), ffma (FFMA persistent kernel),
three (three kernels per env-step), msg (SelectionAction / UpdateQ batches), noalloc (qp_td_grad_noalloc)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

forms = sys.argv[1:] or ["c2", "c2s", "tc", "ffma", "three", "msg"]
OPTS = {"c2": {}, "c2s": {}, "tc": {"ep_c2": 0}, "ffma": {"ep_tc": 0}, "three": {"fused": 0}}
n, steps = 130, 3
for f in forms:
    if f == "msg":
        with q.QDecisionPolicy(q.PolicyConfig()) as pol:
            rng = np.random.default_rng(0)
            x = rng.uniform(10, 20, (70, 203)).astype(np.float32)
            a = pol.select_action(x, 900.0)
            pol.update_q(x, rng.normal(size=70).astype(np.float32), x[::-1].copy())
            print("msg ok", a[:5])
        continue
    prices = O.gen_prices(42, 0, n, 260)
    if f == "c2s":
        prices = prices * np.float32(1.0371)
    with q.QDecisionPolicy(q.PolicyConfig()) as pol:
        for k, v in OPTS[f].items():
            pol.set_option(k, v)
        pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 1)
        act = {k: int(pol.get_option(k)) for k in ("fused_active", "ep_tc_active", "ep_c2_active", "ep_c2_split_active")}
        pol.env_step(steps)
        pol.env_step(steps)
        tr = pol.env_step_trace(2)
        print(f, "ok", act, "steps", pol.steps_done, "portfolio mean", float(np.mean(pol.portfolios())))
