"""GPU tests of the batching front-end (BASELINE.json configs[4]: router fan-out, actors batching
select_action / update_q into one launch) and of the policy checkpoint."""
import threading

import numpy as np
import pytest

from oracle import oracle as O
from tests import _tol
from tests.test_gpu_parity import make_policy, q_mod, states_of

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(201, 200), (62, 256)])
def test_enqueue_flush_equals_direct_batch(shape):
    """Eight 'child actors' on eight threads enqueue their messages; one flush answers them all.  (62, 256): a policy on
    the GEMM-organised path (W1 outside the per-tile kernels' shared memory), where the message-level calls run G1 / row
    kernel / G2."""
    q = q_mod()
    cfg = O.Config(wn=shape[0], hidden=shape[1], seed=3)
    n_threads, per = 8, 40
    n = n_threads * per
    rng = np.random.default_rng(0)
    w = (O.gen_normal(4, 0, cfg.P) * 0.05).astype(np.float32)
    s = rng.uniform(50, 150, (n, cfg.D)).astype(np.float32)
    s2 = s + rng.uniform(-1, 1, (n, cfg.D)).astype(np.float32)
    r = rng.uniform(-5, 5, n).astype(np.float32)
    steps = rng.integers(0, 1200, n).astype(np.float32)
    with make_policy(cfg, params=w) as direct, make_policy(cfg, params=w) as batched:
        assert direct.get_option("wide") == (1 if shape[1] > 240 else 0)
        a_ref = direct.select_action(s, steps, np.arange(n))
        direct.update_q(s, r, s2)
        front = q.BatchingPolicy(batched)
        tickets = [None] * n

        def child(t):
            for j in range(t * per, (t + 1) * per):
                tickets[j] = front.enqueue_select(s[j], float(steps[j]), j)

        th = [threading.Thread(target=child, args=(t,)) for t in range(n_threads)]
        [x.start() for x in th]
        [x.join() for x in th]
        assert front.pending == n and sorted(tickets) == list(range(n))
        with pytest.raises(q.QPolicyError, match="still queued"):
            front.action(tickets[0])
        assert front.flush() == (n, 0)
        got = np.array([int(front.action(tickets[j])) for j in range(n)])
        np.testing.assert_array_equal(got, a_ref)          # same kernel, same rows -> identical
        # updates: the order rows arrive in does not change the set summed, only the fp32 sum order
        order = rng.permutation(n)
        for j in order:
            front.enqueue_update(s[j], float(r[j]), s2[j])
        assert front.flush() == (0, n)
        _tol.assert_grad_close(batched.last_grad(), direct.last_grad())
        assert np.abs(batched.get_weights()[0] - direct.get_weights()[0]).max() <= _tol.W_ATOL
        # guards apply at enqueue time, as in receive (QDecisionPolicyActor.scala:56-57, :64-65)
        with pytest.raises(q.IllegalArgumentException):
            front.enqueue_select(s[0][:47], 0.0, 0)
        with pytest.raises(q.IllegalArgumentException):
            front.enqueue_update(s[0], 1.0, s2[0][:10])
        assert front.pending == 0 and front.flush() == (0, 0)


def test_front_end_queues_grow_and_alternate():
    """The queues are two generations of pinned staging buffers that grow by doubling while other threads keep
    reserving slots: 3 200 selects from eight threads (1 024 -> 4 096 slots, twice: once per generation), tickets
    stay unique and consecutive across flushes, every action is the one a direct batch gives, and a ticket of an
    older flush is refused."""
    q = q_mod()
    cfg = O.Config(seed=5)
    n_threads, per = 8, 400
    n = n_threads * per
    rng = np.random.default_rng(1)
    w = (O.gen_normal(6, 0, cfg.P) * 0.05).astype(np.float32)
    with make_policy(cfg, params=w) as direct, make_policy(cfg, params=w) as batched:
        front = q.BatchingPolicy(batched)
        first_round_ticket = None
        for rnd in range(2):
            s = rng.uniform(50, 150, (n, cfg.D)).astype(np.float32)
            steps = rng.integers(0, 1200, n).astype(np.float32)
            a_ref = direct.select_action(s, steps, np.arange(n))
            tickets = [None] * n

            def child(t):
                for j in range(t * per, (t + 1) * per):
                    tickets[j] = front.enqueue_select(s[j], float(steps[j]), j)

            th = [threading.Thread(target=child, args=(t,)) for t in range(n_threads)]
            [x.start() for x in th]
            [x.join() for x in th]
            assert front.pending == n
            assert sorted(tickets) == list(range(rnd * n, (rnd + 1) * n))
            assert front.flush() == (n, 0)
            got = np.array([int(front.action(tickets[j])) for j in range(n)])
            np.testing.assert_array_equal(got, a_ref)
            if rnd == 0:
                first_round_ticket = tickets[0]
        with pytest.raises(q.IllegalArgumentException, match="most recent flush"):
            front.action(first_round_ticket)
        # updates through the grown queues: one gradient-SUM step over all of them
        s2 = s + rng.uniform(-1, 1, (n, cfg.D)).astype(np.float32)
        r = rng.uniform(-5, 5, n).astype(np.float32)
        direct.update_q(s, r, s2)

        def child_u(t):
            for j in range(t * per, (t + 1) * per):
                front.enqueue_update(s[j], float(r[j]), s2[j])

        th = [threading.Thread(target=child_u, args=(t,)) for t in range(n_threads)]
        [x.start() for x in th]
        [x.join() for x in th]
        assert front.flush() == (0, n)
        _tol.assert_grad_close(batched.last_grad(), direct.last_grad())


def test_router_fanout_in_lockstep():
    """Tickers x episodes through the front-end, stepping in lock-step: equals the fused episode path."""
    q = q_mod()
    cfg = O.Config(wn=16, hidden=20, bookkeeping=O.RLPY, lr=1e-3, seed=5)
    tickers, per = 6, 5
    n = tickers * per
    prices = O.gen_prices(cfg.seed, 0, tickers, 50)
    soe = np.repeat(np.arange(tickers), per).astype(np.int32)
    w = (O.gen_normal(8, 0, cfg.P) * 0.2).astype(np.float32)
    with make_policy(cfg, params=w) as fused, make_policy(cfg, params=w) as pol:
        fused.env_create(prices, soe, 300.0, 1)
        tr = fused.env_step_trace(12)
        front = q.BatchingPolicy(pol)
        budget, shares, sv = np.full(n, 300.0), np.ones(n, np.int32), np.zeros(n)
        for i in range(12):
            s = states_of(cfg, prices, soe, i, budget, shares)
            tk = [front.enqueue_select(s[e], float(i), e) for e in range(n)]
            front.flush()
            a = np.array([int(front.action(t)) for t in tk])
            np.testing.assert_array_equal(a, tr["action_sel"][i])
            p_new = prices[soe, i + cfg.wn].astype(np.float64)
            cur = budget + shares * sv
            nb, ns = budget.copy(), shares.copy()
            buy, sell = (a == 0) & (budget >= p_new), (a == 1) & (shares > 0)
            nb[buy] -= p_new[buy]; ns[buy] += 1
            nb[sell] += p_new[sell]; ns[sell] -= 1
            s2 = states_of(cfg, prices, soe, i + 1, nb, ns)
            rew = (nb + ns * p_new - cur).astype(np.float32)
            for e in range(n):
                front.enqueue_update(s[e], float(rew[e]), s2[e])
            front.flush()
            budget, shares, sv = nb, ns, p_new
            np.testing.assert_array_equal(shares, tr["shares"][i])
        assert np.abs(fused.get_weights()[0] - pol.get_weights()[0]).max() <= 12 * _tol.W_ATOL


def test_checkpoint_round_trip(tmp_path):
    cfg = O.Config(bookkeeping=O.RLPY, lr=1e-4)
    prices = O.gen_prices(cfg.seed, 0, 40, 300)
    soe = np.arange(40, dtype=np.int32)
    w = (O.gen_normal(1, 0, cfg.P) * 0.05).astype(np.float32)
    path = tmp_path / "policy.ckpt"
    with make_policy(cfg, params=w) as a:
        a.env_create(prices, soe, 1000.0, 1)
        a.env_step(20)
        a.save_checkpoint(path)
        a.env_step(15)
        wa, sa, pa = a.get_weights(), a.env_state(), a.portfolios()
    with make_policy(cfg) as b:                      # different (default) weights until the load
        b.env_create(prices, soe, 1000.0, 1)
        b.load_checkpoint(path)
        assert b.steps_done == 20
        b.env_step(15)
        wb, sb, pb = b.get_weights(), b.env_state(), b.portfolios()
    for x, y in zip(wa + sa, wb + sb):
        np.testing.assert_array_equal(x, y)          # resumed run is bit-identical
    np.testing.assert_array_equal(pa, pb)
    q = q_mod()
    with make_policy(O.Config(wn=8, hidden=16)) as c:
        with pytest.raises(q.QPolicyError, match="shape"):
            c.load_checkpoint(path)
        with pytest.raises(q.QPolicyError, match="cannot open"):
            c.load_checkpoint(tmp_path / "missing.ckpt")


def test_snapshot_every_n_updates(tmp_path):
    """QDecisionPolicyActor.scala:74: `if (iteration % 500 == 0 && iteration != 0) saveSnapshot(...)`."""
    cfg = O.Config(wn=8, hidden=16)
    rng = np.random.default_rng(0)
    s = rng.uniform(1, 2, (3, cfg.D)).astype(np.float32)
    path = tmp_path / "snap.ckpt"
    with make_policy(cfg) as pol:
        pol.set_snapshot(path, every=4)
        for i in range(4):
            pol.update_q(s, 1.0, s)
        assert not path.exists()                     # iterations 0..3 seen so far: no snapshot at 0
        pol.update_q(s, 1.0, s)                      # handled with iteration == 4 -> snapshot of the state after it
        assert path.exists()
        w_now = pol.get_weights()[0]
        with make_policy(cfg) as other:
            other.load_checkpoint(path)
            np.testing.assert_array_equal(other.get_weights()[0], w_now)
            assert other.iteration == 5 == pol.iteration     # the snapshot's weights contain five updates
            path.unlink()
            other.set_snapshot(path, every=4)
            other.update_q(s, 1.0, s)                        # handled with iteration 5: no repeat of the snapshot
            assert not path.exists() and other.iteration == 6
