"""Regenerates the golden fixtures in this directory from the CPU oracle (oracle/qp_oracle.c).

The reference cannot run here (no JVM) and its tests assert no numbers, so these
vectors pin (1) the INPUT fixtures the reference's tests supply and the results derived from its
code by hand (final portfolio 2000.0 for TrainerChildActorSpec.scala:59-76), and (2) the oracle's
own outputs, so that the oracle and the CUDA path are both held to the same committed numbers.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def grid_weights(seed, n):
    """N(0,1) draws snapped to a 1/1024 grid: exact in fp32 and immune to libm last-bit differences."""
    return (np.round(O.gen_normal(seed, 0, n).astype(np.float64) * 1024.0) / 1024.0).astype(np.float32)


def trace_arrays(traces):
    keys = ("action_sel", "action_eff", "explored", "k_next", "reward", "q_s", "q_n", "target", "budget", "shares")
    return {k: np.stack([t[k] for t in traces]) for k in keys}


def policy_spec():
    # QDecisionPolicyActorSpec.scala:16-19, :35
    cfg = O.Config()
    s = np.array(list(range(55, 256)) + [1000, 0], np.float32)
    s2 = np.array(list(range(55, 256)) + [940, 1], np.float32)
    w = grid_weights(42, cfg.P)
    o = O.Oracle(cfg, params=w)
    a0, q0, e0 = o.select_action(s, 0.0)
    a5, q5, e5 = o.select_action(s, 500.0)
    g, info = o.td_grad(s, 10.0, s2)
    o.update_q(s, 10.0, s2)
    a_after, q_after, _ = o.select_action(s, 500.0)
    idx = np.linspace(0, cfg.P - 1, 64).astype(np.int64)
    np.savez(os.path.join(HERE, "policy_spec.npz"), state=s, next_state=s2, reward=np.float32(10.0),
             weight_seed=42, weight_checksum=np.int64(np.round(w.astype(np.float64) * 1024).sum()),
             action_step0=a0, explored_step0=e0, action_step500=a5, explored_step500=e5, q_s=info["q_s"],
             q_n=info["q_n"], k_next=info["k_next"], target=info["target"], grad_abs_sum=np.float64(np.abs(g.astype(np.float64)).sum()),
             grad_nonzero=np.int64((g != 0).sum()), sample_idx=idx, grad_sample=g[idx],
             params_after_sample=o.params[idx], acc_after_sample=o.acc[idx], q_after=q_after, action_after=a_after)


def episode(name, cfg, n_series, T, soe, B0, N0, n_steps, weight_scale=1.0, start_step=0):
    prices = O.gen_prices(cfg.seed, 0, n_series, T)
    w = grid_weights(cfg.seed, cfg.P) * np.float32(weight_scale)
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, soe, B0, N0)
    o.env["step"] = start_step
    tr = trace_arrays(o.env_step(n_steps, trace=True))
    idx = np.linspace(0, cfg.P - 1, 64).astype(np.int64)
    np.savez(os.path.join(HERE, name), window=cfg.wn, hidden=cfg.hidden, lr=np.float32(cfg.lr), bookkeeping=cfg.bookkeeping,
             seed=cfg.seed, n_series=n_series, T=T, series_of_env=np.asarray(soe, np.int32), B0=B0, N0=N0,
             n_steps=n_steps, start_step=start_step, weight_scale=np.float32(weight_scale), prices_head=prices[:, :8], sample_idx=idx,
             params_after_sample=o.params[idx], acc_after_sample=o.acc[idx], portfolios=o.portfolios(), **tr)


def trainer_child_spec():
    # TrainerChildActorSpec.scala:59-76: prices 55..256, budget 2000, 0 shares -> one step, portfolio 2000.0
    cfg = O.Config()
    prices = np.arange(55, 257, dtype=np.float32)[None, :]
    o = O.Oracle(cfg, params=grid_weights(42, cfg.P))
    o.env_create(prices, [0], 2000.0, 0)
    tr = trace_arrays(o.env_step(o.n_steps_total, trace=True))
    np.savez(os.path.join(HERE, "trainer_child_spec.npz"), prices=prices, B0=2000.0, N0=0,
             n_steps=o.n_steps_total, portfolio=o.portfolios(), **tr)


if __name__ == "__main__":
    policy_spec()
    trainer_child_spec()
    episode("episode_rlpy.npz", O.Config(bookkeeping=O.RLPY, lr=1e-4), 4, 300, [0, 0, 1, 1, 2, 2, 3, 3], 1000.0, 2, 40)
    episode("episode_late.npz", O.Config(bookkeeping=O.RLPY, lr=1e-4), 6, 1000, [0, 1, 2, 3, 4, 5, 5, 4, 3], 1000.0, 1, 40,
            weight_scale=0.05, start_step=700)
    episode("episode_reference.npz", O.Config(bookkeeping=O.REFERENCE), 3, 260, [0, 1, 2, 0, 1], 1000.0, 3, 30)
    episode("episode_tiny.npz", O.Config(wn=8, hidden=16, bookkeeping=O.RLPY, lr=1e-3), 5, 64, [0, 1, 2, 3, 4, 4, 3], 500.0, 1, 25,
            weight_scale=0.25)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
