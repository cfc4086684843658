"""CPU tests of the oracle itself: golden vectors, known-answer tests, an independent autograd
statement of the reference graph, and the fp64 twin that calibrates the fp32 tolerance."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests import _tol
from tests.golden.make_golden import grid_weights


# Random123 known-answer vectors for Philox4x32-10
PHILOX_KAT = [
    ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
    ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
    ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
     [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
]


@pytest.mark.parametrize("ctr,key,want", PHILOX_KAT)
def test_philox_known_answers(ctr, key, want):
    assert O.philox4x32_10(ctr, key) == want


def test_price_walk_is_quarter_tick_and_reproducible():
    p = O.gen_prices(42, 0, 16, 1000)
    assert p.shape == (16, 1000) and (p[:, 0] == 100.0).all()
    steps = np.diff(p.astype(np.float64), axis=1)
    assert set(np.unique(steps)).issubset({-0.5, -0.25, 0.0, 0.25, 0.5})
    assert (p >= 1.0).all()
    # a shard generates exactly the rows of the whole
    assert np.array_equal(O.gen_prices(42, 5, 3, 1000), p[5:8])


def test_golden_policy_spec(golden_dir):
    """QDecisionPolicyActorSpec.scala:16-19, :35 inputs; outputs pinned from the oracle."""
    g = np.load(os.path.join(golden_dir, "policy_spec.npz"))
    cfg = O.Config()
    assert g["state"].size == 203 and g["next_state"].size == 203  # the spec's own assertion (:23-24)
    w = grid_weights(int(g["weight_seed"]), cfg.P)
    assert int(np.round(w.astype(np.float64) * 1024).sum()) == int(g["weight_checksum"])
    o = O.Oracle(cfg, params=w)
    a0, _, e0 = o.select_action(g["state"], 0.0)
    assert e0[0] == 1 and a0[0] == g["action_step0"][0]           # step 0 always explores (:58)
    a5, _, e5 = o.select_action(g["state"], 500.0)
    assert a5[0] == g["action_step500"][0] and e5[0] == g["explored_step500"][0]
    grad, info = o.td_grad(g["state"], float(g["reward"]), g["next_state"])
    np.testing.assert_array_equal(info["q_s"], g["q_s"])
    np.testing.assert_array_equal(info["q_n"], g["q_n"])
    np.testing.assert_array_equal(info["k_next"], g["k_next"])
    np.testing.assert_array_equal(info["target"], g["target"])
    np.testing.assert_array_equal(grad[g["sample_idx"]], g["grad_sample"])
    assert (grad != 0).sum() == int(g["grad_nonzero"])
    o.update_q(g["state"], float(g["reward"]), g["next_state"])
    np.testing.assert_array_equal(o.params[g["sample_idx"]], g["params_after_sample"])
    np.testing.assert_array_equal(o.acc[g["sample_idx"]], g["acc_after_sample"])
    _, q_after, _ = o.select_action(g["state"], 500.0)
    np.testing.assert_array_equal(q_after, g["q_after"])


def _run_golden_episode(g):
    cfg = O.Config(wn=int(g["window"]), hidden=int(g["hidden"]), lr=float(g["lr"]),
                   bookkeeping=int(g["bookkeeping"]), seed=int(g["seed"]))
    prices = O.gen_prices(cfg.seed, 0, int(g["n_series"]), int(g["T"]))
    np.testing.assert_array_equal(prices[:, :8], g["prices_head"])
    w = grid_weights(cfg.seed, cfg.P) * np.float32(g["weight_scale"])
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, g["series_of_env"], float(g["B0"]), int(g["N0"]))
    o.env["step"] = int(g["start_step"])
    return cfg, o, o.env_step(int(g["n_steps"]), trace=True)


@pytest.mark.parametrize("name", ["episode_rlpy", "episode_reference", "episode_tiny", "episode_late"])
def test_golden_episodes(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    _, o, tr = _run_golden_episode(g)
    for k in ("action_sel", "action_eff", "explored", "k_next", "shares", "budget", "reward", "q_s", "q_n", "target"):
        np.testing.assert_array_equal(np.stack([t[k] for t in tr]), g[k], err_msg=k)
    np.testing.assert_array_equal(o.portfolios(), g["portfolios"])
    np.testing.assert_array_equal(o.params[g["sample_idx"]], g["params_after_sample"])


def test_trainer_child_spec_golden(golden_dir):
    """TrainerChildActorSpec.scala:59-76: 202 prices 55..256, budget 2000, no shares: exactly one
    step; with no shares Sell is demoted to Hold and the portfolio stays 2000.0 (derived)."""
    g = np.load(os.path.join(golden_dir, "trainer_child_spec.npz"))
    assert g["prices"].shape == (1, 202) and int(g["n_steps"]) == 1
    assert g["portfolio"][0] == 2000.0
    cfg = O.Config()
    o = O.Oracle(cfg, params=grid_weights(42, cfg.P))
    o.env_create(g["prices"], [0], 2000.0, 0)
    tr = o.env_step(1, trace=True)[0]
    assert tr["reward"][0] == 0.0 or tr["action_eff"][0] == 0   # Buy costs nothing in value either
    assert o.portfolios()[0] == 2000.0


def test_series_too_short_is_illegal_argument():
    """TrainerChildActorSpec.scala:28-37 (45 prices) and the guard at TrainerChildActor.scala:65, :70."""
    o = O.Oracle(O.Config())
    with pytest.raises(ValueError):
        o.env_create(np.arange(55, 100, dtype=np.float32)[None, :], [0], 2000.0, 0)
    with pytest.raises(ValueError):   # exactly window prices is still too short (size > h1Dim + 1)
        o.env_create(np.arange(201, dtype=np.float32)[None, :], [0], 2000.0, 0)


def test_router_statistics_golden():
    """TrainerRouterActorSpec.scala:75-78: ten portfolios of 10.0 -> avg 10.0, std 0.0; population std."""
    assert O.portfolio_stats(np.full(10, 10.0)) == (10.0, 0.0)
    m, s = O.portfolio_stats(np.array([1.0, 2.0, 3.0, 4.0]))
    assert m == 2.5 and abs(s - np.sqrt(1.25)) < 1e-15


def test_reference_mode_with_no_shares_never_moves_the_portfolio():
    """with N0 = 0 the REFERENCE book-keeping yields reward == 0.0 exactly."""
    cfg = O.Config(bookkeeping=O.REFERENCE)
    p = O.gen_prices(7, 0, 6, 400)
    o = O.Oracle(cfg)
    o.env_create(p, np.arange(6), 2400.0, 0)
    tr = o.env_step(o.n_steps_total, trace=True)
    assert all((t["reward"] == 0.0).all() for t in tr)
    assert (o.portfolios() == 2400.0).all()
    assert set(np.unique(np.stack([t["shares"] for t in tr]))) <= {0, 1}


@pytest.mark.skipif(not os.path.exists("/root/reference/src/main/resources/MSFT-stock-prices-revised.txt"),
                    reason="reference data file only exists in the build container")
def test_msft_series_reference_mode():
    """The reference's own App run (ShareTradeHelper.scala:20-21: budget 2400, 0 shares) on its bundled
    price file: 6047 rows -> 5846 steps; REFERENCE book-keeping leaves every portfolio at 2400.0."""
    rows = []
    with open("/root/reference/src/main/resources/MSFT-stock-prices-revised.txt") as f:
        for line in f:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) >= 2:
                try:
                    rows.append((parts[1], float(parts[0])))
                except ValueError:
                    pass
    rows.sort()
    prices = np.asarray([r[1] for r in rows], np.float64).astype(np.float32)
    assert prices.size == 6047
    o = O.Oracle(O.Config())
    o.env_create(prices[None, :], [0, 0], 2400.0, 0)
    assert o.n_steps_total == 5846
    o.env_step(300)
    assert (o.portfolios() == 2400.0).all()


def test_sequential_equals_batched_for_one_episode():
    """B = 1: the build's gradient-SUM rule is exactly the reference's per-message update."""
    cfg = O.Config(bookkeeping=O.RLPY)
    p = O.gen_prices(3, 0, 1, 260)
    a, b = O.Oracle(cfg), O.Oracle(cfg)
    for o in (a, b):
        o.env_create(p, [0], 1000.0, 1)
    a.env_step(20, mode=0)
    b.env_step(20, mode=1)
    np.testing.assert_array_equal(a.params, b.params)
    np.testing.assert_array_equal(a.acc, b.acc)


def test_multithreaded_baseline_matches_serial_within_tolerance():
    cfg = O.Config(bookkeeping=O.RLPY)
    p = O.gen_prices(5, 0, 24, 230)
    a, b = O.Oracle(cfg), O.Oracle(cfg)
    for o in (a, b):
        o.env_create(p, np.arange(24), 1000.0, 1)
    a.env_step(3)
    b.env_step(3, threads=4)
    np.testing.assert_array_equal(a.env["shares"], b.env["shares"])
    np.testing.assert_array_equal(a.env["budget"], b.env["budget"])
    assert np.abs(a.params - b.params).max() <= _tol.W_ATOL


def test_autograd_statement_of_the_reference_graph():
    """Independent check of the hand-derived backward and of ResourceApplyAdagrad: the graph of
    QDecisionPolicyActor.scala:38-50 written with torch autograd (CPU, fp64)."""
    torch = pytest.importorskip("torch")
    cfg = O.Config(wn=12, hidden=16)
    rng = np.random.default_rng(1)
    B = 7
    w = (rng.standard_normal(cfg.P) * 0.3).astype(np.float32)
    s = rng.uniform(0.5, 3.0, (B, cfg.D)).astype(np.float32)
    s2 = rng.uniform(0.5, 3.0, (B, cfg.D)).astype(np.float32)
    r = rng.uniform(-1, 1, B).astype(np.float32)
    o = O.Oracle(cfg, np.float64, params=w)
    g, info = o.td_grad(s, r, s2)
    assert (info["q_s"] > 0).any(), "fixture must exercise live output units"

    W1 = torch.tensor(w[: cfg.D * cfg.hidden].reshape(cfg.D, cfg.hidden), dtype=torch.float64, requires_grad=True)
    W2 = torch.tensor(w[cfg.D * cfg.hidden:].reshape(cfg.hidden, 3), dtype=torch.float64, requires_grad=True)

    bias = float(np.float32(cfg.bias))   # the constants are fp32 in the reference (:42, :46)

    def net(x):
        h1 = torch.relu(x @ W1 + bias)
        return torch.relu(h1 @ W2 + bias)

    xs, xn = torch.tensor(s, dtype=torch.float64), torch.tensor(s2, dtype=torch.float64)
    with torch.no_grad():
        q_s, q_n = net(xs), net(xn)
        k = q_n.argmax(dim=1)
        y = q_s.clone()
        y[torch.arange(B), k] = torch.tensor(r, dtype=torch.float64) + np.float64(np.float32(cfg.gamma)) * q_n[torch.arange(B), k]
    loss = torch.square(y - net(xs))      # un-reduced, as at :49; minimize() sums it
    loss.sum().backward()
    g_ref = np.concatenate([W1.grad.numpy().ravel(), W2.grad.numpy().ravel()])
    np.testing.assert_allclose(g, g_ref, rtol=1e-10, atol=1e-12)
    np.testing.assert_array_equal(info["k_next"], k.numpy())
    # AdaGrad: accum += g^2; var -= lr * g / sqrt(accum)
    acc0 = o.acc.copy()
    o.update_q(s, r, s2)
    acc_ref = acc0 + g_ref ** 2
    np.testing.assert_allclose(o.acc, acc_ref, rtol=1e-12)
    np.testing.assert_allclose(o.params, w.astype(np.float64) - np.float64(np.float32(cfg.lr)) * g_ref / np.sqrt(acc_ref), rtol=1e-10, atol=1e-12)


def test_fp32_vs_fp64_twin():
    """Calibration of tests/_tol.py: the fp32 oracle against its fp64 twin, teacher-forced per step."""
    cfg = O.Config(bookkeeping=O.RLPY)
    p = O.gen_prices(42, 0, 32, 1000)
    rng = np.random.default_rng(0)
    worst = 0.0
    for trial in range(6):
        w = O.gen_normal(100 + trial, 0, cfg.P)
        o32, o64 = O.Oracle(cfg, np.float32, params=w), O.Oracle(cfg, np.float64, params=w)
        i = int(rng.integers(0, 799))
        for o in (o32, o64):
            o.env_create(p, np.arange(32), 1000.0, 2)
            o.env["step"] = i
        x = np.concatenate([p[:, i:i + cfg.wn], np.full((32, 1), 1000.0, np.float32), np.full((32, 1), 2.0, np.float32)], axis=1)
        t32 = o32.env_step(1, trace=True, want_grad=True)[0]
        t64 = o64.env_step(1, trace=True, want_grad=True)[0]
        worst = max(worst, _tol.assert_q_close(cfg, w, x, t32["q_s"], t64["q_s"]))
        _tol.assert_actions_match(cfg, w, x, t32["action_sel"], t64["action_sel"], t64["q_s"])
        np.testing.assert_array_equal(t32["shares"], t64["shares"])
        _tol.assert_grad_close(t32["grad"], t64["grad"])
        assert np.abs(o32.params - o64.params).max() <= _tol.W_ATOL
    assert worst < 0.5, f"fp32 oracle uses {worst:.2f} of the tolerance; it should sit well inside"


def test_teacher_hook_imposes_choices_and_nothing_else():
    """oracle.env_step(force_action=, force_knext=): entries < 0 leave the oracle's own choice; imposing the
    oracle's own choices reproduces the unforced run bit for bit; imposing another action changes exactly that
    episode's decision (TrainerChildActor.scala:118-123) and the TD index enters the gradient column."""
    cfg = O.Config(wn=8, hidden=16, bookkeeping=O.RLPY, seed=3)
    prices = O.gen_prices(cfg.seed, 0, 6, 40)
    soe = np.arange(6, dtype=np.int32)
    w = (np.abs(O.gen_normal(1, 0, cfg.P)) * 0.05).astype(np.float32)      # all-positive: every ReLU alive

    def run(fa=None, fk=None):
        o = O.Oracle(cfg, params=w)
        o.env_create(prices, soe, 500.0, 1)
        o.env["step"] = 5
        return o.env_step(1, trace=True, want_grad=True, force_action=fa, force_knext=fk)[0], o

    ref, o0 = run()
    same, o1 = run(ref["action_sel"], ref["k_next"])
    for k in ref:
        np.testing.assert_array_equal(ref[k], same[k], err_msg=k)
    np.testing.assert_array_equal(o0.params, o1.params)
    none, _ = run(np.full(6, -1, np.int32), np.full(6, -1, np.int32))
    np.testing.assert_array_equal(ref["grad"], none["grad"])
    fa = np.full(6, -1, np.int32)
    fa[2] = (ref["action_sel"][2] + 1) % 3
    alt, _ = run(fa)
    assert alt["action_sel"][2] == fa[2]
    others = np.arange(6) != 2
    np.testing.assert_array_equal(alt["action_sel"][others], ref["action_sel"][others])
    np.testing.assert_array_equal(alt["budget"][others], ref["budget"][others])
    fk = np.full(6, -1, np.int32)
    fk[4] = (ref["k_next"][4] + 1) % 3
    altk, _ = run(None, fk)
    assert altk["k_next"][4] == fk[4] and not np.array_equal(altk["grad"], ref["grad"])
