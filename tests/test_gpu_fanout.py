"""The router fan-out path (csrc/qp_fanout.cuh): many episodes on few price series — the reference's own shape
(TrainerRouterActor.scala:44, 60-66, 88 broadcasts ONE series to all children; BASELINE.json configs[4] is 512 tickers x
256 episodes).  The window part of the hidden layer is computed once per (series, step) and the backward of the window
rows is an outer product per series, so the sums over episodes are re-associated: compared with the oracle at the
stated tolerance, integer / fp64 book-keeping bit for bit.
"""
import numpy as np
import pytest

from oracle import oracle as O
from tests import _tol
from tests._teacher import Tally, teacher_forced_step
from tests.test_gpu_parity import make_policy

pytestmark = pytest.mark.gpu


def fan_policy(cfg, params=None, force=True):
    pol = make_policy(cfg, params=params)
    if force:
        pol.set_option("fanout", 2)
    return pol


# (series, episodes per series): one chunk, several chunks with a ragged last one, many short runs, ragged counts
@pytest.mark.parametrize("bookkeeping", [O.REFERENCE, O.RLPY])
@pytest.mark.parametrize("n_series,per", [(1, 10), (3, 300), (40, 7), (5, 256), (7, None)])
def test_fanout_teacher_forced(bookkeeping, n_series, per):
    cfg = O.Config(bookkeeping=bookkeeping, seed=91)
    counts = [per] * n_series if per else [1, 2, 255, 256, 257, 31, 600]
    soe = np.repeat(np.arange(n_series), counts).astype(np.int32)
    n = soe.size
    prices = O.gen_prices(cfg.seed, 0, n_series, 400) * np.float32(1.0371)     # (any prices: nothing here needs tf32 values)
    w0 = (O.gen_normal(21, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w0)
    o.env_create(prices, soe, 1000.0, 1, env_id_base=10 ** 10)
    # different positions inside a series, so that the budget / shares terms differ between its episodes
    o.env["budget"][:] = 1000.0 - 3.25 * (np.arange(n) % 97)
    o.env["shares"][:] = np.arange(n) % 4
    tally = Tally()
    with fan_policy(cfg) as pol:
        pol.env_create(prices, soe, 1000.0, 1, env_id_base=10 ** 10)
        assert pol.get_option("fanout_active") == 1 and pol.get_option("fused_active") == 0
        for i in (0, 1, 57, 150, 198):
            teacher_forced_step(cfg, o, pol, prices, soe, i, f"fan-out {n_series}x{per} step {i}", tally)
    tally.assert_mostly_unforced()


def test_fanout_config5_shape_against_oracle():
    """series_of_env = repeat(tickers, 256), unscaled N(0,1) weights (the reference's initialisation), 64 tickers:
    16 384 episodes, teacher-forced at two steps; gradient sums over that many episodes are arbitrated by the fp64 twin
    (tests/_teacher.py)."""
    cfg = O.Config(bookkeeping=O.RLPY, seed=42)
    tickers, per = 64, 256
    soe = np.repeat(np.arange(tickers), per).astype(np.int32)
    n = soe.size
    prices = O.gen_prices(42, 0, tickers, 600)
    o = O.Oracle(cfg)
    o.env_create(prices, soe, 1000.0, 0)
    tally = Tally()
    with make_policy(cfg) as pol:                       # default options: the fan-out path is chosen on its own
        pol.env_create(prices, soe, 1000.0, 0)
        assert pol.get_option("fanout_active") == 1
        for i in (0, 300):
            teacher_forced_step(cfg, o, pol, prices, soe, i, f"fan-out 64x256 step {i}", tally, n_terms=n)
    tally.assert_mostly_unforced()


def test_fanout_is_chosen_only_where_it_pays():
    cfg = O.Config()
    prices = O.gen_prices(1, 0, 4096, 260)
    with make_policy(cfg) as pol:
        pol.env_create(prices, np.arange(4096, dtype=np.int32), 1000.0, 0)          # one episode per series
        assert pol.get_option("fanout_active") == 0 and pol.get_option("ep_c2_active") == 1
        pol.env_create(prices[:16], np.repeat(np.arange(16), 256).astype(np.int32), 1000.0, 0)   # fits the pair kernel
        assert pol.get_option("fanout_active") == 0 and pol.get_option("ep_c2_active") == 1
        pol.env_create(prices[:32], np.repeat(np.arange(32), 256).astype(np.int32), 1000.0, 0)   # 8 192 episodes
        assert pol.get_option("fanout_active") == 1 and pol.get_option("fused_active") == 0
        soe = np.repeat(np.arange(32), 256).astype(np.int32)[::-1].copy()           # not grouped in ascending order
        pol.env_create(prices[:32], soe, 1000.0, 0)
        assert pol.get_option("fanout_active") == 0
        pol.set_option("fanout", 0)
        pol.env_create(prices[:32], np.repeat(np.arange(32), 256).astype(np.int32), 1000.0, 0)
        assert pol.get_option("fanout_active") == 0


def test_fanout_free_running_matches_three_kernel_path():
    """The same 20 series x 30 episodes for 40 env-steps on the fan-out path and on the three-kernel path: identical
    book-keeping up to the first action that differs (there is none at this size), weights within the re-association."""
    cfg = O.Config(bookkeeping=O.RLPY, lr=1e-4, seed=9)
    soe = np.repeat(np.arange(20), 30).astype(np.int32)
    prices = O.gen_prices(cfg.seed, 0, 20, 400)
    w = (O.gen_normal(3, 0, cfg.P) * 0.05).astype(np.float32)
    res = []
    for fan in (True, False):
        pol = fan_policy(cfg, params=w) if fan else make_policy(cfg, params=w, fused=0)
        with pol:
            if not fan:
                pol.set_option("fanout", 0)
            pol.env_create(prices, soe, 1000.0, 1)
            assert pol.get_option("fanout_active") == (1 if fan else 0)
            pol.env_set_state(step=120)
            tr = pol.env_step_trace(25)
            pol.env_step(15)
            res.append((tr, pol.get_weights()[0], pol.env_state(), pol.steps_done))
    (ta, wa, sa, da), (tb, wb, sb, db) = res
    assert da == db == 160
    for k in ("action_sel", "action_eff", "explored", "shares", "budget", "reward", "k_next"):
        np.testing.assert_array_equal(ta[k], tb[k], err_msg=k)
    np.testing.assert_allclose(ta["q_s"], tb["q_s"], rtol=1e-4, atol=1e-3)
    assert np.abs(wa - wb).max() <= 40 * _tol.W_ATOL
    for x, y in zip(sa, sb):
        np.testing.assert_array_equal(x, y)


def test_fanout_bit_reproducible():
    cfg = O.Config(bookkeeping=O.RLPY, seed=5)
    soe = np.repeat(np.arange(12), 700).astype(np.int32)
    prices = O.gen_prices(cfg.seed, 0, 12, 300)
    outs = []
    for _ in range(2):
        with make_policy(cfg) as pol:
            pol.env_create(prices, soe, 1000.0, 0)
            assert pol.get_option("fanout_active") == 1
            pol.env_step(30)
            outs.append((pol.get_weights()[0], pol.portfolios()))
    np.testing.assert_array_equal(outs[0][0], outs[1][0])
    np.testing.assert_array_equal(outs[0][1], outs[1][1])
