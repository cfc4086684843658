"""Parity of the CUDA path against the CPU oracle, through the C ABI (ctypes -> libqpolicy.so).

Bit-exact: action indices, explore flags, share counts, budgets, rewards (integer and fp64
book-keeping).  Within the stated fp32 tolerance (tests/_tol.py): Q-values, targets, gradients,
updated weights; an action may differ only where the oracle's top-two Q gap is inside tolerance.
"""
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests import _tol
from tests._teacher import Tally, states_of, teacher_forced_step
from tests.golden.make_golden import grid_weights

pytestmark = pytest.mark.gpu


def q_mod():
    import scala_akka_tensorflow_sharetrade_helper_b200 as q
    return q


# name -> (fused, ep_tc, ep_c2).  "c2s" is the pair kernel on prices that are NOT tf32-exact (32 real + 32 virtual episodes
# per pair: x = tf32(x) + remainder, both through the same MMAs); the test multiplies its prices by INEXACT for it.
PATHS = {"c2": (1, 1, 1), "c2s": (1, 1, 1), "tc": (1, 1, 0), "ffma": (1, 0, 0), "three": (0, 0, 0)}
INEXACT = np.float32(1.0371)


def prices_for(path, prices):
    return prices * INEXACT if path == "c2s" else prices


def make_policy(cfg: O.Config, params=None, acc=None, fused=None, ep_tc=None, path=None):
    """path: "c2" = persistent tcgen05 kernel on CTA pairs that split the hidden dimension (the default where it
    applies), "tc" = persistent tcgen05 kernel, one tile per CTA, "ffma" = persistent FFMA kernel, "three" = three
    kernels per env-step.  fused / ep_tc set the two options one by one; `fused` alone keeps its old meaning
    (the FFMA form of the persistent kernel / the three-kernel path)."""
    q = q_mod()
    pc = q.PolicyConfig(window=cfg.wn, hidden=cfg.hidden, epsilon=cfg.epsilon, gamma=cfg.gamma, lr=cfg.lr,
                        bias=cfg.bias, step_div=cfg.step_div, adagrad_init_acc=cfg.adagrad_init_acc,
                        bookkeeping=cfg.bookkeeping, seed=cfg.seed)
    pol = q.QDecisionPolicy(pc)
    if params is not None or acc is not None:
        pol.set_weights(params, acc)
    ep_c2 = None
    if path is not None:
        fused, ep_tc, ep_c2 = PATHS[path]
    elif fused is not None and ep_tc is None:
        ep_tc = 0
    if fused is not None:
        pol.set_option("fused", int(fused))
    if ep_tc is not None:
        pol.set_option("ep_tc", int(ep_tc))
    if ep_c2 is not None:
        pol.set_option("ep_c2", int(ep_c2))
    if path is not None or fused is not None:
        pol.set_option("fanout", 0)          # (a test that names its path means that kernel, not the fan-out path)
    return pol


def assert_path(pol, path):
    fused, ep_tc, ep_c2 = PATHS[path]
    got = (pol.get_option("fused_active"), pol.get_option("ep_tc_active"), pol.get_option("ep_c2_active"))
    assert got == (fused, ep_tc, ep_c2), f"path {path} not taken: (fused, ep_tc, ep_c2) active = {got}"
    assert pol.get_option("ep_c2_split_active") == (1 if path == "c2s" else 0), f"path {path}: wrong form of the pair kernel"


# ------------------------------------------------------------------------------------------------
# construction, weights, errors
# ------------------------------------------------------------------------------------------------
def test_default_init_matches_oracle_generator():
    """qp_create draws W ~ N(0,1) from the seed exactly as the oracle's host generator does."""
    cfg = O.Config(seed=1234)
    with make_policy(cfg) as pol:
        p, a = pol.get_weights()
    np.testing.assert_array_equal(p, O.gen_normal(1234, 0, cfg.P))
    assert (a == np.float32(1e-8)).all()
    assert abs(float(p.std()) - 1.0) < 0.02


def test_size_guards_raise_illegal_argument():
    """QDecisionPolicyActorSpec.scala:47-65: a 47-float state is rejected for both messages."""
    q = q_mod()
    good = np.array(list(range(55, 256)) + [1000, 0], np.float32)
    wrong = np.array(list(range(55, 100)) + [1000, 0], np.float32)
    with make_policy(O.Config()) as pol:
        with pytest.raises(q.IllegalArgumentException, match="size"):
            pol.receive(q.SelectionAction(wrong, 0))
        with pytest.raises(q.IllegalArgumentException, match="size"):
            pol.receive(q.UpdateQ(good, 10, wrong))
        with pytest.raises(q.IllegalArgumentException):
            pol.receive(q.UpdateQ(wrong, 10, good))
        # TrainerChildActorSpec.scala:28-37: 45 prices
        with pytest.raises(q.IllegalArgumentException, match="Stock price count"):
            pol.env_create(np.arange(55, 100, dtype=np.float32)[None, :], [0], 2000.0, 0)
        with pytest.raises(q.QPolicyError):
            pol.env_step(1)        # no episodes yet
        # still usable afterwards
        assert isinstance(pol.receive(q.SelectionAction(good, 0)), q.Action)


def test_unsupported_shape_is_reported():
    q = q_mod()
    with pytest.raises(q.QPolicyError, match="shape"):
        make_policy(O.Config(wn=513, hidden=2052))     # fits neither the per-tile path nor the GEMM path


# ------------------------------------------------------------------------------------------------
# message level: SelectionAction / UpdateQ
# ------------------------------------------------------------------------------------------------
def test_policy_spec_fixture(golden_dir):
    """The reference's own spec inputs (QDecisionPolicyActorSpec.scala:16-19, :35) through the
    message API, against the committed golden outputs of the oracle."""
    q = q_mod()
    g = np.load(os.path.join(golden_dir, "policy_spec.npz"))
    cfg = O.Config()
    w = grid_weights(int(g["weight_seed"]), cfg.P)
    s, s2 = g["state"], g["next_state"]
    with make_policy(cfg, params=w) as pol:
        a0 = pol.receive(q.SelectionAction(s, 0))                     # "return action in initial stage"
        assert isinstance(a0, q.Action) and int(a0) == int(g["action_step0"][0])
        act, qv = pol.select_action(s[None, :], 500.0, return_q=True)
        _tol.assert_q_close(cfg, w, s[None, :], qv, g["q_s"])
        assert int(act[0]) == int(g["action_step500"][0])
        assert isinstance(pol.receive(q.UpdateQ(s, 10, s2)), q.Updated)  # "Updated"
        grad = pol.last_grad()
        assert (grad != 0).sum() == int(g["grad_nonzero"])
        np.testing.assert_allclose(grad[g["sample_idx"]], g["grad_sample"], rtol=1e-4, atol=1e-4 * np.abs(g["grad_sample"]).max())
        p, a = pol.get_weights()
        assert np.abs(p[g["sample_idx"]] - g["params_after_sample"]).max() <= _tol.W_ATOL
        act2, q2 = pol.select_action(s[None, :], 500.0, return_q=True)   # "return an action after updateQ"
        _tol.assert_q_close(cfg, p, s[None, :], q2, g["q_after"])
        assert int(act2[0]) == int(g["action_after"][0])


@pytest.mark.parametrize("B", [1, 7, 32, 33, 257])
@pytest.mark.parametrize("shape", [(201, 200), (8, 16), (30, 44)])
def test_select_action_batches(B, shape):
    cfg = O.Config(wn=shape[0], hidden=shape[1], seed=11)
    rng = np.random.default_rng(B * 1000 + shape[0])
    w = (O.gen_normal(5, 0, cfg.P) * 0.1).astype(np.float32)
    x = rng.uniform(50, 150, (B, cfg.D)).astype(np.float32)
    steps = rng.integers(0, 1200, B).astype(np.float32)
    steps[0] = 0.0
    ids = rng.integers(0, 2 ** 40, B).astype(np.int64)
    o = O.Oracle(cfg, params=w)
    a_ref, q_ref, e_ref = o.select_action(x, steps, ids)
    with make_policy(cfg, params=w) as pol:
        a, qv = pol.select_action(x, steps, ids, return_q=True)
        a_noid = pol.select_action(x, steps)
    _tol.assert_q_close(cfg, w, x, qv, q_ref)
    _tol.assert_actions_match(cfg, w, x, a, a_ref, q_ref)
    # explored rows depend on Philox alone: bit-exact
    np.testing.assert_array_equal(a[e_ref == 1], a_ref[e_ref == 1])
    a_ref2, _, _ = o.select_action(x, steps)
    _tol.assert_actions_match(cfg, w, x, a_noid, a_ref2, q_ref)


def test_argmax_ties_go_to_buy():
    """Dead hidden units make the three Q-values equal (relu(0 + 0.1)); Tensor.argmax returns the
    first maximal index -> Buy."""
    cfg = O.Config(wn=8, hidden=16)
    w = -np.abs(O.gen_normal(3, 0, cfg.P)) - 1.0          # all-negative weights: h == 0, q == bias everywhere
    x = np.random.default_rng(0).uniform(1, 2, (40, cfg.D)).astype(np.float32)
    with make_policy(cfg, params=w) as pol:
        a, qv = pol.select_action(x, 5000.0, np.arange(40), return_q=True)   # thr = 0.9
    _, _, e_ref = O.Oracle(cfg, params=w).select_action(x, 5000.0, np.arange(40))
    assert (qv == np.float32(0.1)).all()
    assert (a[e_ref == 0] == 0).all() and (e_ref == 0).sum() > 20


@pytest.mark.parametrize("B", [1, 5, 32, 100])
@pytest.mark.parametrize("shape", [(201, 200), (8, 16)])
def test_update_q_batches(B, shape):
    cfg = O.Config(wn=shape[0], hidden=shape[1])
    rng = np.random.default_rng(B + shape[1])
    w = (O.gen_normal(9, 0, cfg.P) * (0.05 if shape[0] > 100 else 0.3)).astype(np.float32)
    acc = rng.uniform(0.5, 2.0, cfg.P).astype(np.float32)
    s = rng.uniform(50, 150, (B, cfg.D)).astype(np.float32)
    s2 = s + rng.uniform(-1, 1, (B, cfg.D)).astype(np.float32)
    r = rng.uniform(-5, 5, B).astype(np.float32)
    o = O.Oracle(cfg, params=w, acc=acc)
    g_ref, info = o.td_grad(s, r, s2)
    assert np.abs(g_ref).max() > 0, "fixture must produce a live gradient"
    o.update_q(s, r, s2)
    with make_policy(cfg, params=w, acc=acc) as pol:
        pol.update_q(s, r, s2)
        g = pol.last_grad()
        p, a = pol.get_weights()
    _tol.assert_grad_close(g, g_ref)
    # acc = acc0 + g^2: twice the relative error of g, which inherits the Q tolerance through dq = 2(q - target)
    np.testing.assert_allclose(a, o.acc, rtol=1e-3)
    assert np.abs(p - o.params).max() <= _tol.W_ATOL


# ------------------------------------------------------------------------------------------------
# episode level
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("path", ["c2", "c2s", "tc", "ffma", "three"])
@pytest.mark.parametrize("bookkeeping", [O.REFERENCE, O.RLPY])
@pytest.mark.parametrize("n_envs,N0", [(1, 0), (5, 3), (32, 1), (33, 0), (64, 1), (65, 1), (200, 2)])
def test_env_step_teacher_forced(bookkeeping, n_envs, N0, path):
    """Each step starts from the oracle's weights and episode state; device and oracle take one step;
    everything the step produces is compared (tests/_teacher.py), on each of the three execution paths."""
    cfg = O.Config(bookkeeping=bookkeeping, seed=77)
    n_series = max(1, n_envs // 2)
    prices = prices_for(path, O.gen_prices(cfg.seed, 0, n_series, 400))    # quarter ticks: tf32-exact (c2s: times 1.0371, not exact)
    soe = (np.arange(n_envs) % n_series).astype(np.int32)
    w0 = (O.gen_normal(21, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w0)
    o.env_create(prices, soe, 1000.0, N0, env_id_base=10 ** 10)
    tally = Tally()
    with make_policy(cfg, path=path) as pol:
        pol.env_create(prices, soe, 1000.0, N0, env_id_base=10 ** 10)
        assert_path(pol, path)
        for i in (0, 1, 57, 150, 198):
            teacher_forced_step(cfg, o, pol, prices, soe, i, f"{path} step {i}", tally)
    tally.assert_mostly_unforced()
    if path == "tc":     # the one-tile-per-CTA kernel needs tf32-exact prices: without the pair kernel the FFMA form serves them
        with make_policy(cfg, path="tc") as pol:
            pol.env_create(prices * INEXACT, soe, 1000.0, N0)
            assert pol.get_option("ep_tc_active") == 0 and pol.get_option("fused_active") == 1


GOLDEN_CASES = [(n, p_) for n in ("episode_rlpy", "episode_reference", "episode_late") for p_ in ("c2", "tc", "ffma", "three")] + \
               [("episode_tiny", "three")]     # (window 8 / hidden 16 is outside the persistent kernels' shapes)


@pytest.mark.parametrize("name,path", GOLDEN_CASES)
def test_golden_episodes_free_running(golden_dir, name, path):
    """Committed golden traces (oracle-generated): the device runs the same episodes free-running on each
    execution path; integer book-keeping must match bit-for-bit up to the first tolerated action divergence.
    episode_late starts at env-step 700 (exploit threshold 0.7), so most of its actions are argmax Q(s)."""
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    cfg = O.Config(wn=int(g["window"]), hidden=int(g["hidden"]), lr=float(g["lr"]),
                   bookkeeping=int(g["bookkeeping"]), seed=int(g["seed"]))
    prices = O.gen_prices(cfg.seed, 0, int(g["n_series"]), int(g["T"]))
    w = grid_weights(cfg.seed, cfg.P) * np.float32(g["weight_scale"])
    n_steps, i0 = int(g["n_steps"]), int(g["start_step"])
    with make_policy(cfg, params=w, path=path) as pol:
        pol.env_create(prices, g["series_of_env"], float(g["B0"]), int(g["N0"]))
        assert_path(pol, path)
        if i0:
            pol.env_set_state(step=i0)
        tr = pol.env_step_trace(n_steps)
        ports = pol.portfolios()
    diverged = tr["action_sel"] != g["action_sel"]
    first_div = np.where(diverged.any(axis=1))[0]
    upto = int(first_div[0]) if first_div.size else n_steps
    assert upto >= min(n_steps, 10), f"free-running parity lost already at step {upto}"
    for k in ("action_sel", "action_eff", "explored", "k_next", "shares", "budget", "reward"):
        np.testing.assert_array_equal(tr[k][:upto], g[k][:upto], err_msg=k)
    scale = np.maximum(1.0, np.abs(g["q_s"][:upto]).max())
    assert np.abs(tr["q_s"][:upto] - g["q_s"][:upto]).max() <= 1e-3 * scale   # weights drift apart slowly when free-running
    if upto == n_steps:
        np.testing.assert_array_equal(ports, g["portfolios"])


@pytest.mark.parametrize("path", ["c2", "c2s", "tc"])
def test_tensor_core_kernel_free_running_exploit_branch(path):
    """The tensor-core persistent kernel free-running against the oracle with the reference's ε schedule in its
    exploit regime (steps 600..629: 60-63 % of the actions are argmax Q(s), QDecisionPolicyActor.scala:58-60), four
    tiles of episodes, one launch.  Integer book-keeping bit for bit up to the first action divergence, which must
    be a tolerated one (oracle's top-two gap inside the Q tolerance at the weights of that step)."""
    cfg = O.Config(bookkeeping=O.RLPY, lr=1e-4, seed=17)
    n, T, i0, steps = 128, 900, 600, 30
    prices = prices_for(path, O.gen_prices(cfg.seed, 0, n, T))
    soe = np.arange(n, dtype=np.int32)
    w = (O.gen_normal(6, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, soe, 1000.0, 1)
    o.env["step"] = i0
    ref, w_at, st_at = [], [], []
    for _ in range(steps):
        w_at.append(o.params.copy())
        st_at.append((o.env["budget"].copy(), o.env["shares"].copy()))
        ref.append(o.env_step(1, trace=True)[0])
    with make_policy(cfg, params=w, path=path) as pol:
        pol.env_create(prices, soe, 1000.0, 1)
        assert_path(pol, path)
        pol.env_set_state(step=i0)
        tr = pol.env_step_trace(steps)
    R = {k: np.stack([r[k] for r in ref]) for k in ("action_sel", "action_eff", "explored", "k_next", "shares", "budget", "reward", "q_s")}
    assert (R["explored"] == 0).mean() > 0.5, "fixture must exercise the exploit branch"
    diverged = (tr["action_sel"] != R["action_sel"]) | (tr["k_next"] != R["k_next"])
    first = np.where(diverged.any(axis=1))[0]
    upto = int(first[0]) if first.size else steps
    assert upto >= 10, f"free-running parity lost already at step {upto}"
    for k in ("action_sel", "action_eff", "explored", "k_next", "shares", "budget", "reward"):
        np.testing.assert_array_equal(tr[k][:upto], R[k][:upto], err_msg=k)
    assert (np.abs(tr["q_s"][:upto] - R["q_s"][:upto]) <= 1e-3 * np.maximum(1.0, np.abs(R["q_s"][:upto]).max())).all()
    if upto < steps and (tr["action_sel"][upto] != R["action_sel"][upto]).any():
        x = states_of(cfg, prices, soe, i0 + upto, *st_at[upto])
        _tol.assert_actions_match(cfg, w_at[upto], x, tr["action_sel"][upto], R["action_sel"][upto], R["q_s"][upto],
                                  f"first divergence (step {i0 + upto})")


@pytest.mark.parametrize("init_acc", [1e-8, 0.1])
@pytest.mark.parametrize("path", ["c2", "tc", "three"])
def test_adagrad_initial_accumulator_candidates(init_acc, path):
    """QDecisionPolicyActor.scala:50 `AdaGrad(0.01f)`: the initial accumulator is tensorflow_scala's default, taken
    on trust as 1e-8; TF-Python's 0.1 is the other candidate.  It is a config field: both values go through
    qp_create (not set_weights) and one full step is compared with the oracle."""
    cfg = O.Config(bookkeeping=O.RLPY, seed=23, adagrad_init_acc=init_acc)
    n = 96
    prices = O.gen_prices(cfg.seed, 0, n, 300)
    soe = np.arange(n, dtype=np.int32)
    w = (O.gen_normal(4, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w)
    assert (o.acc == np.float32(init_acc)).all()
    o.env_create(prices, soe, 1000.0, 1)
    with make_policy(cfg, path=path) as pol:
        assert (pol.get_weights()[1] == np.float32(init_acc)).all()          # what qp_create installed
        pol.set_weights(w, None)
        pol.env_create(prices, soe, 1000.0, 1)
        assert_path(pol, path)
        pol.env_step(1)
        o.env_step(1)
        p1, a1 = pol.get_weights()
        g = pol.last_grad()
        np.testing.assert_allclose(a1, o.acc, rtol=1e-3)
        live = np.abs(g) > 1e-3 * np.abs(g).max()
        assert live.sum() > 1000
        # the step size is what tells the two candidates apart: lr*sign(g) from 1e-8, lr*g/sqrt(0.1+g^2) from 0.1
        assert np.abs(p1 - o.params)[live].max() <= _tol.W_ATOL
        tally = Tally()
        teacher_forced_step(cfg, o, pol, prices, soe, 1, f"{path} acc0={init_acc}", tally)
        tally.assert_mostly_unforced()


def test_trainer_child_spec_through_the_actor_mirror(golden_dir):
    """TrainerChildActorSpec.scala:59-76 through TrainerChild.receive(Train): Trained, then 2000.0."""
    q = q_mod()
    with make_policy(O.Config()) as pol:
        t = q.TrainerChild(pol, 2000, 0)
        assert t.receive("GetPortfolio") == q.NotComputed()
        data = q.StockDataResponse("my-share", {f"d{i:04d}": float(i) for i in range(55, 257)})
        assert t.receive("Train", data) == q.Trained()
        assert t.receive("GetPortfolio") == q.TrainedData(2000.0)
        assert t.receive("Initialise") == q.Initialised()
        assert t.receive("GetPortfolio") == q.NotComputed()


def test_message_api_reproduces_the_fused_episode_path():
    """Driving SelectionAction / UpdateQ from a host loop that restates TrainerChildActor.scala:87-102
    gives the same trajectory as the fused env path (same kernels' arithmetic, same Philox counters)."""
    cfg = O.Config(wn=16, hidden=20, bookkeeping=O.RLPY, lr=1e-3, seed=5)
    prices = O.gen_prices(cfg.seed, 0, 3, 60)
    soe = np.array([0, 1, 2, 1], np.int32)
    w = (O.gen_normal(8, 0, cfg.P) * 0.2).astype(np.float32)
    with make_policy(cfg, params=w) as fused, make_policy(cfg, params=w) as msg:
        fused.env_create(prices, soe, 300.0, 1)
        tr = fused.env_step_trace(30)
        budget = np.full(4, 300.0)
        shares = np.ones(4, np.int32)
        sv = np.zeros(4)
        for i in range(30):
            s = states_of(cfg, prices, soe, i, budget, shares)
            cur = budget + shares * sv
            a = msg.select_action(s, float(i), np.arange(4))
            np.testing.assert_array_equal(a, tr["action_sel"][i])
            p_new = prices[soe, i + cfg.wn].astype(np.float64)
            nb, ns = budget.copy(), shares.copy()
            buy = (a == 0) & (budget >= p_new)
            sell = (a == 1) & (shares > 0)
            nb[buy] -= p_new[buy]; ns[buy] += 1
            nb[sell] += p_new[sell]; ns[sell] -= 1
            reward = (nb + ns * p_new - cur).astype(np.float32)
            s2 = states_of(cfg, prices, soe, i + 1, nb, ns)
            msg.update_q(s, reward, s2)
            budget, shares, sv = nb, ns, p_new
            np.testing.assert_array_equal(shares, tr["shares"][i])
            np.testing.assert_array_equal(budget, tr["budget"][i])
        pf, _ = fused.get_weights()
        pm, _ = msg.get_weights()
        assert np.abs(pf - pm).max() <= 1e-6


def test_persistent_kernel_matches_three_kernel_path():
    """The same 300 episodes for 40 env-steps through the persistent cooperative kernel and through
    the three-kernels-per-step path: identical book-keeping, weights equal within the gradient-sum
    re-association (the two paths split the sum over episodes differently)."""
    cfg = O.Config(bookkeeping=O.RLPY, lr=1e-4, seed=9)
    n = 300
    prices = O.gen_prices(cfg.seed, 0, 50, 400)
    soe = (np.arange(n) % 50).astype(np.int32)
    w = (O.gen_normal(3, 0, cfg.P) * 0.05).astype(np.float32)
    res = []
    for fused in (1, 0):
        with make_policy(cfg, params=w, fused=fused) as pol:
            pol.env_create(prices, soe, 1000.0, 1)
            pol.env_set_state(step=120)
            assert pol.get_option("fused_active") == fused
            tr = pol.env_step_trace(25)
            pol.env_step(15)                     # continue without trace: state carried across launches
            res.append((tr, pol.get_weights()[0], pol.env_state(), pol.steps_done))
    (ta, wa, sa, da), (tb, wb, sb, db) = res
    assert da == db == 160
    for k in ("action_sel", "action_eff", "explored", "shares", "budget", "reward", "k_next"):
        np.testing.assert_array_equal(ta[k], tb[k], err_msg=k)
    np.testing.assert_allclose(ta["q_s"], tb["q_s"], rtol=1e-4, atol=1e-3)
    for x, y in zip(sa, sb):
        np.testing.assert_array_equal(x, y)
    assert np.abs(wa - wb).max() <= 40 * _tol.W_ATOL


@pytest.mark.parametrize("path", ["c2", "tc"])
def test_persistent_tensor_core_kernel_free_running(path):
    """40 env-steps in one launch (weights republished as tf32 hi/lo every step inside the kernel),
    against the FFMA persistent kernel on the same episodes."""
    cfg = O.Config(bookkeeping=O.RLPY, lr=1e-4, seed=9)
    n = 300
    prices = O.gen_prices(cfg.seed, 0, 50, 400)
    soe = (np.arange(n) % 50).astype(np.int32)
    w = (O.gen_normal(3, 0, cfg.P) * 0.05).astype(np.float32)
    res = []
    for pth in (path, "ffma"):
        with make_policy(cfg, params=w, path=pth) as pol:
            pol.env_create(prices, soe, 1000.0, 1)
            pol.env_set_state(step=120)
            assert_path(pol, pth)
            tr = pol.env_step_trace(25)
            pol.env_step(15)
            res.append((tr, pol.get_weights()[0], pol.env_state(), pol.steps_done))
    (ta, wa, sa, da), (tb, wb, sb, db) = res
    assert da == db == 160
    for k in ("action_sel", "action_eff", "explored", "shares", "budget", "reward", "k_next"):
        np.testing.assert_array_equal(ta[k], tb[k], err_msg=k)
    np.testing.assert_allclose(ta["q_s"], tb["q_s"], rtol=1e-4, atol=1e-3)
    assert np.abs(wa - wb).max() <= 40 * _tol.W_ATOL


@pytest.mark.parametrize("shape", [(200, 192), (220, 180)])
def test_persistent_tensor_core_kernel_other_shapes(shape):
    """The edges of the shapes the tensor-core kernel accepts: hidden = 192 (the smallest N whose operand ring
    holds the epilogue's staging boxes; Np == H, no padding columns) and window 220 / hidden 180 (12 padding
    columns in the MMA's N, the last 32-column box of the partial clipped by the tensor map at column 180),
    teacher-forced against the oracle."""
    cfg = O.Config(wn=shape[0], hidden=shape[1], bookkeeping=O.RLPY, seed=5)
    n_envs = 70
    prices = O.gen_prices(cfg.seed, 0, 35, cfg.wn + 60)
    soe = (np.arange(n_envs) % 35).astype(np.int32)
    w0 = (O.gen_normal(8, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w0)
    o.env_create(prices, soe, 1000.0, 1)
    tally = Tally()
    with make_policy(cfg) as pol:
        pol.env_create(prices, soe, 1000.0, 1)
        assert pol.get_option("ep_tc_active") == 1, "shape not taken by the tensor-core kernel"
        for i in (0, 31, 58):
            teacher_forced_step(cfg, o, pol, prices, soe, i, f"ep_tc {shape} step {i}", tally)
    tally.assert_mostly_unforced()


@pytest.mark.parametrize("path", ["c2", "c2s", "tc"])
def test_persistent_tensor_core_kernel_ring_wraps(path):
    """330 env-steps in ONE launch.  tc: the window ring (256 rows) wraps more than once, every row block of the
    dW1 partial takes the wrapped-store path in turn; c2: the window buffer is copied back to its top six times.  step_div is huge, so the ε-greedy threshold stays 0 and
    every action is the Philox draw (independent of Q): the book-keeping trajectory is then identical for the
    tensor-core kernel, the FFMA kernel and the oracle, and Q / weights can be compared all the way."""
    cfg = O.Config(bookkeeping=O.RLPY, lr=1e-5, seed=31, step_div=1e12)      # (lr small enough to keep the ReLUs alive)
    n, T, steps = 96, 201 + 340, 330
    prices = prices_for(path, O.gen_prices(cfg.seed, 0, n, T))
    soe = np.arange(n, dtype=np.int32)
    w = (O.gen_normal(4, 0, cfg.P) * 0.05).astype(np.float32)
    res = []
    for pth in (path, "ffma"):
        with make_policy(cfg, params=w, path=pth) as pol:
            pol.env_create(prices, soe, 1000.0, 2)
            assert_path(pol, pth)
            tr = pol.env_step_trace(steps)
            res.append((tr, pol.get_weights()[0], pol.last_grad()))
    (ta, wa, ga), (tb, wb, gb) = res
    assert ta["explored"].all()
    for k in ("action_sel", "action_eff", "shares", "budget", "reward"):
        np.testing.assert_array_equal(ta[k], tb[k], err_msg=k)
    scale = np.maximum(1.0, np.abs(tb["q_s"]).max(axis=2, keepdims=True))
    assert (np.abs(ta["q_s"] - tb["q_s"]) <= 2e-3 * scale).all(), "Q(s) of the two persistent kernels drifted apart"
    assert np.abs(wa - wb).max() <= steps * _tol.W_ATOL
    d = np.linalg.norm(ga - gb) / np.linalg.norm(gb)
    assert d <= 2e-3, f"last gradient: relative Frobenius distance {d:.3g}"
    # and the oracle, over the whole stretch
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, soe, 1000.0, 2)
    ref = o.env_step(steps, trace=True)
    for k in ("action_sel", "action_eff", "shares", "budget", "reward"):
        np.testing.assert_array_equal(ta[k], np.stack([r[k] for r in ref]), err_msg=f"oracle {k}")
    q_ref = np.stack([r["q_s"] for r in ref])
    scale = np.maximum(1.0, np.abs(q_ref).max(axis=2, keepdims=True))
    assert (np.abs(ta["q_s"] - q_ref) <= 5e-3 * scale).all(), "Q(s) drifted away from the oracle"
    assert np.abs(wa - o.params).max() <= steps * _tol.W_ATOL


def test_wave_mode_is_bit_reproducible():
    """The tensor-core kernel in waves (several tiles per CTA): the later waves ADD their dW1 partial to the first
    wave's through the memory system (TMA reduce-add).  Only the CTA that owns a partial adds to it, one wave
    after the other, so the sum has a fixed order: two runs give the same bits."""
    cfg = O.Config(bookkeeping=O.RLPY, seed=3)
    n = 15000
    prices = O.gen_prices(cfg.seed, 0, 300, 260)
    soe = (np.arange(n) % 300).astype(np.int32)
    w = (O.gen_normal(5, 0, cfg.P) * 0.05).astype(np.float32)
    res = []
    for _ in range(2):
        with make_policy(cfg, params=w) as pol:
            pol.env_create(prices, soe, 1000.0, 1)
            assert pol.get_option("ep_tc_active") == 1
            pol.env_step(6)
            res.append((pol.get_weights()[0], pol.last_grad(), pol.env_state()))
    np.testing.assert_array_equal(res[0][0], res[1][0])
    np.testing.assert_array_equal(res[0][1], res[1][1])
    for a_, b_ in zip(res[0][2], res[1][2]):
        np.testing.assert_array_equal(a_, b_)
    assert np.abs(res[0][1]).max() > 0


# ------------------------------------------------------------------------------------------------
# GEMM-organised path (BASELINE.json configs[3]: hist 512 -> window 513, hidden 2048)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape,n_envs", [((40, 256), 70), ((30, 128), 300), ((513, 2048), 130)])
@pytest.mark.parametrize("bookkeeping", [O.REFERENCE, O.RLPY])
def test_wide_path_teacher_forced(shape, n_envs, bookkeeping):
    cfg = O.Config(wn=shape[0], hidden=shape[1], bookkeeping=bookkeeping, seed=13)
    T = shape[0] + 40
    n_series = max(1, n_envs // 3)
    prices = O.gen_prices(cfg.seed, 0, n_series, T)
    soe = (np.arange(n_envs) % n_series).astype(np.int32)
    w0 = (O.gen_normal(2, 0, cfg.P) * (0.02 if shape[1] > 1000 else 0.05)).astype(np.float32)
    o = O.Oracle(cfg, params=w0)
    o.env_create(prices, soe, 1000.0, 2, env_id_base=77)
    tally = Tally()
    with make_policy(cfg) as pol:
        pol.set_option("wide", 1)
        assert pol.get_option("wide") == 1 and pol.get_option("fused_active") == 0
        pol.env_create(prices, soe, 1000.0, 2, env_id_base=77)
        for i in (0, 17, 38):
            teacher_forced_step(cfg, o, pol, prices, soe, i, f"wide step {i}", tally)
    tally.assert_mostly_unforced()


@pytest.mark.parametrize("shape,n_envs,T,wscale", [((40, 256), 70, 80, 0.05), ((130, 128), 200, 172, 0.05), ((513, 2048), 130, 556, 0.02),
                                                   ((513, 2048), 130, 556, 1.0)])
@pytest.mark.parametrize("bookkeeping", [O.REFERENCE, O.RLPY])
def test_wide_path_tensor_cores_teacher_forced(shape, n_envs, T, wscale, bookkeeping):
    """The same step with the two GEMMs on tcgen05 tensor cores (3xTF32: tf32 hi/lo operand pairs,
    three MMAs per k-block).  Held to the SAME fp32 tolerance as the FFMA path.  wscale = 1.0: the reference's
    own N(0,1) initialisation (QDecisionPolicyActor.scala:41, :45), |Q| ~ 1e5 at this width."""
    cfg = O.Config(wn=shape[0], hidden=shape[1], bookkeeping=bookkeeping, seed=21)
    prices = O.gen_prices(cfg.seed, 0, n_envs, T) * np.float32(1.0371)   # not tf32-exact: exercises the lo operands
    soe = np.arange(n_envs, dtype=np.int32)                               # episode e <-> series e (TMA reads the price matrix)
    w0 = (O.gen_normal(2, 0, cfg.P) * np.float32(wscale)).astype(np.float32)
    o = O.Oracle(cfg, params=w0)
    o.env_create(prices, soe, 1000.0, 2)
    tally = Tally()
    with make_policy(cfg) as pol:
        pol.set_option("wide", 1)
        pol.env_create(prices, soe, 1000.0, 2)
        assert pol.get_option("tc_active") == 1
        for i in (0, 11, T - shape[0] - 1):
            teacher_forced_step(cfg, o, pol, prices, soe, i, f"tc step {i}", tally)
        # the FFMA GEMMs on the same state agree with the tensor-core ones within the same tolerance
        pol.set_option("tc", 0)
        assert pol.get_option("tc_active") == 0
        teacher_forced_step(cfg, o, pol, prices, soe, 5, "ffma gemm step 5", tally)
    tally.assert_mostly_unforced()


def test_wide_path_matches_tile_path():
    """A shape both paths can run (window 30, hidden 128): same episodes, 20 steps, free-running."""
    cfg = O.Config(wn=30, hidden=128, bookkeeping=O.RLPY, lr=1e-4, seed=4)
    n = 200
    prices = O.gen_prices(cfg.seed, 0, 40, 120)
    soe = (np.arange(n) % 40).astype(np.int32)
    w = (O.gen_normal(6, 0, cfg.P) * 0.1).astype(np.float32)
    res = []
    for wide in (1, 0):
        with make_policy(cfg, params=w, fused=0) as pol:
            pol.set_option("wide", wide)
            pol.env_create(prices, soe, 500.0, 1)
            pol.env_set_state(step=40)
            tr = pol.env_step_trace(20)
            res.append((tr, pol.get_weights()[0], pol.env_state()))
    (ta, wa, sa), (tb, wb, sb) = res
    for k in ("action_sel", "action_eff", "explored", "shares", "budget", "reward", "k_next"):
        np.testing.assert_array_equal(ta[k], tb[k], err_msg=k)
    np.testing.assert_allclose(ta["q_s"], tb["q_s"], rtol=1e-4, atol=1e-3)
    assert np.abs(wa - wb).max() <= 20 * _tol.W_ATOL


@pytest.mark.parametrize("B", [1, 70, 300])
@pytest.mark.parametrize("shape,wscale", [((513, 2048), 0.02), ((30, 128), 0.1)])
def test_wide_policy_message_level_calls(shape, wscale, B):
    """SelectionAction / UpdateQ (QDecisionPolicyActor.scala:56-76) on the GEMM-organised path — BASELINE.json
    configs[3]'s 515 -> 2048 policy, whose W1 does not fit in shared memory — against the oracle, as
    test_select_action_batches / test_update_q_batches do for the reference shape; and through the batching front-end."""
    cfg = O.Config(wn=shape[0], hidden=shape[1], seed=13)
    rng = np.random.default_rng(B + shape[1] + 1000)
    w = (O.gen_normal(5, 0, cfg.P) * np.float32(wscale)).astype(np.float32)
    acc = rng.uniform(0.5, 2.0, cfg.P).astype(np.float32)
    x = rng.uniform(50, 150, (B, cfg.D)).astype(np.float32)
    x2 = x + rng.uniform(-1, 1, (B, cfg.D)).astype(np.float32)
    r = rng.uniform(-5, 5, B).astype(np.float32)
    steps = rng.integers(0, 1200, B).astype(np.float32)
    ids = rng.integers(0, 2 ** 40, B).astype(np.int64)
    o = O.Oracle(cfg, params=w, acc=acc)
    a_ref, q_ref, e_ref = o.select_action(x, steps, ids)
    g_ref, info = o.td_grad(x, r, x2)
    assert np.abs(g_ref).max() > 0, "fixture must produce a live gradient"
    o.update_q(x, r, x2)
    with make_policy(cfg, params=w, acc=acc) as pol:
        if shape[1] <= 240:
            pol.set_option("wide", 1)
        assert pol.get_option("wide") == 1
        a, qv = pol.select_action(x, steps, ids, return_q=True)
        _tol.assert_q_close(cfg, w, x, qv, q_ref)
        _tol.assert_actions_match(cfg, w, x, a, a_ref, q_ref)
        np.testing.assert_array_equal(a[e_ref == 1], a_ref[e_ref == 1])      # explored rows depend on Philox alone
        with pytest.raises(q_mod().IllegalArgumentException):                # the size guard (:56-57) holds here too
            pol.select_action(x[:, :-1], steps)
        pol.update_q(x, r, x2)
        g = pol.last_grad()
        p, a1 = pol.get_weights()
    _tol.assert_grad_close(g, g_ref)
    # acc = acc0 + g^2 with |dg| <= GRAD_RTOL * max|g| per element (a million parameters: some have a small g next to the largest)
    dg = _tol.GRAD_RTOL * np.abs(g_ref).max()
    assert (np.abs(a1.astype(np.float64) - o.acc) <= 1e-3 * o.acc + 2 * np.abs(g_ref) * dg + dg * dg).all()
    assert np.abs(p - o.params).max() <= _tol.W_ATOL


# ------------------------------------------------------------------------------------------------
# full-size, size-independent properties (BASELINE.json configs[1]: 4096 episodes)
# ------------------------------------------------------------------------------------------------
def test_full_size_determinism_and_conservation():
    cfg = O.Config(bookkeeping=O.REFERENCE)
    n = 4096
    prices = O.gen_prices(42, 0, n, 1000)
    soe = np.arange(n, dtype=np.int32)
    outs = []
    for _ in range(2):
        with make_policy(cfg) as pol:
            pol.env_create(prices, soe, 1000.0, 0)
            assert pol.steps_total == 799
            pol.env_step(799)
            assert pol.steps_done == 799
            with pytest.raises(q_mod().IllegalArgumentException):
                pol.env_step(1)                      # episode finished
            outs.append((pol.portfolios(), pol.get_weights()[0], pol.env_state()[1], pol.portfolio_stats()))
    # bit-reproducible: no atomics, fixed reduction order
    np.testing.assert_array_equal(outs[0][0], outs[1][0])
    np.testing.assert_array_equal(outs[0][1], outs[1][1])
    # Q1: REFERENCE book-keeping from (1000, 0) never changes the portfolio value
    assert (outs[0][0] == 1000.0).all()
    assert set(np.unique(outs[0][2])) <= {0, 1}
    assert outs[0][3] == (1000.0, 0.0)
    assert np.isfinite(outs[0][1]).all()


def test_full_size_rlpy_bookkeeping_invariants():
    cfg = O.Config(bookkeeping=O.RLPY, lr=1e-4)
    n = 4096
    prices = O.gen_prices(42, 0, n, 1000)
    w = (O.gen_normal(42, 0, cfg.P) * 0.05).astype(np.float32)
    with make_policy(cfg, params=w) as pol:
        pol.env_create(prices, np.arange(n, dtype=np.int32), 1000.0, 0)
        tr = pol.env_step_trace(60)
        b, s, v = pol.env_state()
        ports = pol.portfolios()
        mean, std = pol.portfolio_stats()
    # share counts follow the effective actions exactly; budgets never go negative; quarter ticks
    delta = np.where(tr["action_eff"] == 0, 1, np.where(tr["action_eff"] == 1, -1, 0)).sum(axis=0)
    np.testing.assert_array_equal(s, delta)
    assert (s >= 0).all() and (b >= 0).all()
    assert np.array_equal(b * 4, np.round(b * 4))
    np.testing.assert_array_equal(v, prices[:, 60 + cfg.wn - 1].astype(np.float64))
    np.testing.assert_array_equal(ports, b + s * v)
    # the sum of rewards telescopes to the change in portfolio value (fp64 exact on quarter ticks)
    np.testing.assert_array_equal(tr["reward"].astype(np.float64).sum(axis=0), ports - 1000.0)
    assert abs(mean - ports.mean()) < 1e-9 and abs(std - ports.std()) < 1e-9
    # oracle spot-check of the first steps of a slice of episodes is covered by the teacher-forced tests


def test_first_step_of_full_batch_against_oracle():
    """4096 episodes, one step, full comparison with the oracle (seconds on one core)."""
    cfg = O.Config(bookkeeping=O.RLPY)
    n = 4096
    prices = O.gen_prices(42, 0, n, 1000)
    soe = np.arange(n, dtype=np.int32)
    w = (O.gen_normal(42, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, soe, 1000.0, 2)
    tally = Tally()
    with make_policy(cfg, params=w) as pol:
        pol.env_create(prices, soe, 1000.0, 2)
        tr, ref = teacher_forced_step(cfg, o, pol, prices, soe, 640, "full batch", tally, n_terms=n)
    assert (ref["explored"] == 0).sum() > 2000          # exploit path exercised (thr = 0.64)
    tally.assert_mostly_unforced()


@pytest.mark.parametrize("path", ["c2", "c2s", "tc"])
@pytest.mark.parametrize("bookkeeping", [O.REFERENCE, O.RLPY])
def test_bench_regime_unscaled_weights_teacher_forced(bookkeeping, path):
    """bench.py's own numeric regime on the headline kernel: BASELINE.json configs[1] — 4096 episodes, seed-42
    quarter-tick series, and the UNSCALED N(0,1) weights qp_create draws (the reference's initialisation,
    QDecisionPolicyActor.scala:41, :45; |Q| ~ 1e4..1e5), teacher-forced at the first, second, a middle and the
    last env-step of the episode.  The 2xTF32 split (22-bit products) is held to the same tolerance here."""
    cfg = O.Config(bookkeeping=bookkeeping, seed=42)
    n = 2368 if path == "c2s" else 4096                 # (c2s: as many episodes as the pair kernel takes at 32 per pair)
    prices = prices_for(path, O.gen_prices(42, 0, n, 1000))
    soe = np.arange(n, dtype=np.int32)
    o = O.Oracle(cfg)                                   # params = gen_normal(42, 0, P), as qp_create
    o.env_create(prices, soe, 1000.0, 0)
    tally = Tally()
    with make_policy(cfg, path=path) as pol:            # default weights; "c2" is what the default options (bench.py) select
        np.testing.assert_array_equal(pol.get_weights()[0], o.params)
        pol.env_create(prices, soe, 1000.0, 0)
        assert_path(pol, path)
        q_abs = []
        for i in (0, 1, 400, 798):
            tr, ref = teacher_forced_step(cfg, o, pol, prices, soe, i, f"bench regime step {i}", tally, n_terms=n)
            q_abs.append(float(np.abs(ref["q_s"]).max()))
    assert max(q_abs) > 1e3, "fixture is not in the large-magnitude regime"
    tally.assert_mostly_unforced()
    print(f"bench regime {path} bookkeeping {bookkeeping}: max |Q| {max(q_abs):.3g}; gradient distance from the fp64 twin "
          f"(device, fp32 oracle): {tally.grad_err_vs_twin}; imposed choices {tally.forced_actions}+{tally.forced_knext} of {tally.rows}")


@pytest.mark.parametrize("bookkeeping", [O.REFERENCE, O.RLPY])
def test_pair_kernel_in_waves_one_series_per_episode(bookkeeping):
    """BASELINE.json configs[2]'s per-GPU shape in small: every episode on its own series (episode e = series e), more
    tiles than CTA pairs (9 500 episodes = 149 tiles over 74 pairs: 2-3 per pair, the last one ragged).  The pair kernel
    then fetches a tile's window with TMA from the transposed price matrix.  Teacher-forced at two steps, then free-running
    against the one-tile-per-CTA kernel."""
    cfg = O.Config(bookkeeping=bookkeeping, seed=40)
    n = 9500
    prices = O.gen_prices(cfg.seed, 0, n, 280)
    soe = np.arange(n, dtype=np.int32)
    w = (O.gen_normal(15, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, soe, 1000.0, 1)
    tally = Tally()
    with make_policy(cfg, params=w, path="c2") as pol:
        pol.env_create(prices, soe, 1000.0, 1)
        assert_path(pol, "c2")
        for i in (0, 41):
            teacher_forced_step(cfg, o, pol, prices, soe, i, f"pair kernel in waves, step {i}", tally, n_terms=n)
    tally.assert_mostly_unforced()
    res = []
    for path in ("c2", "tc"):
        with make_policy(cfg, params=w * np.float32(0.2), path=path) as pol:      # (small weights, lr as is: ReLUs stay alive)
            pol.env_create(prices, soe, 1000.0, 1)
            tr = pol.env_step_trace(6)
            res.append((tr, pol.get_weights()[0], pol.env_state()))
    (ta, wa, sa), (tb, wb, sb) = res
    for k in ("explored", "reward"):
        np.testing.assert_array_equal(ta[k][0], tb[k][0], err_msg=k)
    same = (ta["action_sel"] == tb["action_sel"]).mean()
    assert same > 0.999, f"the two tensor-core kernels agree on {same:.4f} of the actions only"


@pytest.mark.parametrize("path", ["three", "tc", "c2"])
@pytest.mark.parametrize("n,wscale", [(19000, 0.05), (40007, 0.05), (19000, 1.0)])
def test_large_batch_path_against_oracle(n, wscale, path):
    """More than 128 episodes per SM.  three: the three-kernel path with the large-batch dW1 kernel (one split per
    SM, 16 x 8 register tiles); tc: the tensor-core persistent kernel in waves (5 / 9 tiles per CTA, the last wave
    ragged, the last tile partly filled), whose later waves add their dW1 partial to the first wave's; c2: the pair
    kernel in waves (297 / 626 tiles of 64 episodes over 74 pairs: 4-5 / 8-9 tiles per pair and step, accumulated in TMEM).  One
    teacher-forced step against the oracle (wscale = 1.0: the unscaled N(0,1) weights of the bench), then three more
    free-running steps."""
    cfg = O.Config(bookkeeping=O.RLPY, seed=31)
    prices = O.gen_prices(cfg.seed, 0, 500, 300)
    soe = (np.arange(n) % 500).astype(np.int32)
    w = (O.gen_normal(12, 0, cfg.P) * np.float32(wscale)).astype(np.float32)
    i = 66
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, soe, 1000.0, 1)
    tally = Tally()
    with make_policy(cfg, params=w, path=path) as pol:
        pol.env_create(prices, soe, 1000.0, 1)
        assert_path(pol, path)
        tr, ref = teacher_forced_step(cfg, o, pol, prices, soe, i, f"large batch {path}", tally, n_terms=n)
        pol.env_step(3)
        assert pol.steps_done == i + 4
    tally.assert_mostly_unforced()
