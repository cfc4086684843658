"""SURVEY.md §8 f-3 on the GPU: the reference's own price series (MSFT, 6 047 daily quotes in cents, NOT tf32-exact)
from the raw text through qp_parse_prices into episodes on the device.

The fixture tests/golden/msft_prices.npz carries the bytes of src/main/resources/MSFT-stock-prices-revised.txt
(tests/golden/make_msft_fixture.py), because /root/reference does not exist on the GPU box.
"""
import hashlib
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests import _tol
from tests._teacher import Tally, states_of, teacher_forced_step
from tests.test_gpu_parity import PATHS, assert_path, make_policy, q_mod

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def msft(golden_dir):
    g = np.load(os.path.join(golden_dir, "msft_prices.npz"))
    raw = g["text"].tobytes()
    assert hashlib.sha256(raw).hexdigest() == str(g["sha256"])
    prices = q_mod().parse_prices(raw)              # SharePriceGetter.scala:89-101 + TrainerChildActor.scala:66, in the library
    np.testing.assert_array_equal(prices, g["prices"])
    assert prices.size == 6047 and prices[0] == np.float32(56.080002)
    return prices


def msft_path(pol):
    """The persistent kernel serves these episodes; which form of it is asserted where the test needs it."""
    assert pol.get_option("fused_active") == 1
    if pol.get_option("ep_c2_active") == 1:
        return "c2s" if pol.get_option("ep_c2_split_active") == 1 else "c2"
    return "tc" if pol.get_option("ep_tc_active") == 1 else "ffma"


def test_share_trade_helper_run_reference_bookkeeping(msft):
    """ShareTradeHelper.scala:20-21 + TrainerRouterActor.scala:36, 60-66: budget 2400, no shares, ten children on the one
    series, the reference's N(0,1) initialisation, all 5 846 env-steps (TrainerChildActor.scala:64-71).  With the
    reference's stateless book-keeping every reward is exactly 0.0 and every final portfolio exactly
    2400.0, whatever the policy does: avg / std = (2400.0, 0.0) (TrainerRouterActor.scala:149-151)."""
    q = q_mod()
    cfg = O.Config(bookkeeping=O.REFERENCE)
    soe = np.zeros(10, np.int32)
    with make_policy(cfg) as pol:
        ports = pol.train(msft[None, :], soe, 2400.0, 0)
        assert pol.steps_total == 5846 and pol.steps_done == 5846
        path = msft_path(pol)
        # cent quotes are not tf32 values: the tensor-core pair kernel takes them as tf32(x) + remainder
        assert path == "c2s" and pol.get_option("ep_tc_active") == 1
        assert (ports == 2400.0).all()
        assert pol.portfolio_stats() == (2400.0, 0.0)
        b, s, v = pol.env_state()
        assert set(np.unique(s)) <= {0, 1} and (v == np.float64(msft[-1])).all()
        assert np.isfinite(pol.get_weights()[0]).all()
        # and through the router mirror, traced: rewards identically zero on a stretch of the series
        pol.env_create(msft[None, :], soe, 2400.0, 0)
        tr = pol.env_step_trace(400)
        assert (tr["reward"] == 0.0).all() and (tr["budget"] + tr["shares"] * msft[201:601, None].astype(np.float64) == 2400.0).all()
    with q.QDecisionPolicy(q.PolicyConfig()) as pol2:
        router = q.TrainerRouter(pol2, n_children=10, budget=2400.0, shares=0)
        assert router.start_training() == "NoTrainingDataReceived"
        router.send_training_data(q.StockDataResponse("MSFT", {i: float(p) for i, p in enumerate(msft)}))
        assert router.start_training() == "Completed" and router.is_everything_done() == "Completed"
        assert router.get_avg() == 2400.0 and router.get_std() == 0.0
    print("MSFT run served by the", path, "form of the persistent kernel")


@pytest.mark.parametrize("wscale", [0.05, 1.0])
@pytest.mark.parametrize("path", ["default", "ffma", "three"])
def test_msft_rlpy_teacher_forced(msft, path, wscale):
    """rl.py book-keeping (shares and budget carried) on the real quotes, teacher-forced against the oracle at the
    first steps, mid-series and the last step, on each execution path the series can take; wscale = 1.0 is the
    reference's unscaled initialisation.  Prices like 56.080002 need all 24 mantissa bits."""
    cfg = O.Config(bookkeeping=O.RLPY, seed=7)
    n = 40
    soe = np.zeros(n, np.int32)
    prices = msft[None, :]
    w = (O.gen_normal(3, 0, cfg.P) * np.float32(wscale)).astype(np.float32)
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, soe, 2400.0, 2)
    # give the episodes different positions, so that budget / shares columns differ across the batch
    o.env["budget"][:] = 2400.0 - 37.5 * np.arange(n)
    o.env["shares"][:] = np.arange(n) % 5
    tally = Tally()
    with make_policy(cfg, path=None if path == "default" else path) as pol:
        pol.env_create(prices, soe, 2400.0, 2)
        if path == "default":
            assert msft_path(pol) == "c2s"
        else:
            assert_path(pol, path)
        for i in (0, 1, 2900, 5845):
            teacher_forced_step(cfg, o, pol, prices, soe, i, f"MSFT {path} step {i}", tally)
    tally.assert_mostly_unforced()


def test_msft_rlpy_free_running(msft):
    """300 env-steps from the start of the series in one launch against the oracle run live: integer book-keeping bit
    for bit up to the first action divergence, which must be a tolerated one."""
    cfg = O.Config(bookkeeping=O.RLPY, lr=1e-4, seed=7)
    n, steps = 10, 300
    soe = np.zeros(n, np.int32)
    prices = msft[None, :]
    w = (O.gen_normal(3, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, soe, 2400.0, 0)
    ref, w_at, st_at = [], [], []
    for _ in range(steps):
        w_at.append(o.params.copy())
        st_at.append((o.env["budget"].copy(), o.env["shares"].copy()))
        ref.append(o.env_step(1, trace=True)[0])
    with make_policy(cfg, params=w) as pol:
        pol.env_create(prices, soe, 2400.0, 0)
        assert msft_path(pol) == "c2s"
        tr = pol.env_step_trace(steps)
    R = {k: np.stack([r[k] for r in ref]) for k in ("action_sel", "action_eff", "explored", "k_next", "shares", "budget", "reward", "q_s")}
    diverged = (tr["action_sel"] != R["action_sel"]) | (tr["k_next"] != R["k_next"])
    first = np.where(diverged.any(axis=1))[0]
    upto = int(first[0]) if first.size else steps
    assert upto >= 20, f"free-running parity lost already at step {upto}"
    for k in ("action_sel", "action_eff", "explored", "k_next", "shares", "budget", "reward"):
        np.testing.assert_array_equal(tr[k][:upto], R[k][:upto], err_msg=k)
    if upto < steps and (tr["action_sel"][upto] != R["action_sel"][upto]).any():
        x = states_of(cfg, prices, soe, upto, *st_at[upto])
        _tol.assert_actions_match(cfg, w_at[upto], x, tr["action_sel"][upto], R["action_sel"][upto], R["q_s"][upto],
                                  f"first divergence (step {upto})")
