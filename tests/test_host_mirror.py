"""CPU tests of the host-side mirror of the reference's actor API (no device needed: the policy is
a stand-in that records calls, as the reference's own specs use TestProbe)."""
import numpy as np
import pytest

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from scala_akka_tensorflow_sharetrade_helper_b200.policy import PolicyConfig


class ProbePolicy:
    """Plays the role of `policyProbe` in TrainerChildActorSpec.scala:61-69."""

    def __init__(self):
        self.config = PolicyConfig()
        self.calls = []

    def train(self, prices, series_of_env, budget, shares, env_id_base=0):
        self.calls.append((np.asarray(prices).shape, list(series_of_env), budget, shares))
        return np.full(len(series_of_env), float(budget))

    def portfolio_stats(self):
        return 10.0, 0.0


def stock(n, start=55):
    return q.StockDataResponse("my-share", {f"2001-07-{i:05d}": float(i) for i in range(start, start + n)})


def test_actions_keep_the_reference_order():
    assert [a.name for a in q.Action] == ["Buy", "Sell", "Hold"] and int(q.Hold) == 2


def test_stock_data_iterates_like_a_treemap():
    s = q.StockDataResponse("x", {"2001-01-03": 3.0, "2001-01-01": 1.0, "2001-01-02": 2.0})
    assert s.sorted_prices().tolist() == [1.0, 2.0, 3.0]
    assert s.sorted_prices().dtype == np.float32   # .toFloat at TrainerChildActor.scala:66


def test_trainer_rejects_short_series_like_the_reference():
    """TrainerChildActorSpec.scala:28-37: 45 prices -> IllegalArgumentException."""
    t = q.TrainerChild(ProbePolicy(), 2000, 0)
    with pytest.raises(q.IllegalArgumentException):
        t.receive("Train", stock(45))
    with pytest.raises(q.IllegalArgumentException):
        t.receive("Train", stock(201))


def test_trainer_fsm():
    """TrainerChildActorSpec.scala:39-55."""
    p = ProbePolicy()
    t = q.TrainerChild(p, 2000, 0)
    assert t.receive("GetPortfolio") == q.NotComputed()
    assert t.receive("Train", stock(202)) == q.Trained()
    assert p.calls == [((1, 202), [0], 2000.0, 0)]
    assert t.receive("GetPortfolio") == q.TrainedData(2000.0)
    assert t.receive("Train", stock(202)) is None and len(p.calls) == 1      # "it's already trained"
    assert t.receive("Initialise") == q.Initialised()
    assert t.receive("GetPortfolio") == q.NotComputed()


def test_router_contract():
    """TrainerRouterActorSpec.scala:46-79: no data -> NoTrainingDataReceived; NotComputed before; avg/std after."""
    p = ProbePolicy()
    r = q.TrainerRouter(p, n_children=10, budget=2400.0)
    assert r.start_training() == "NoTrainingDataReceived"
    assert r.get_avg() == q.NotComputed() and r.is_everything_done() == "TrainingNotCompleted"
    r.send_training_data(stock(300))
    assert r.start_training() == "Completed"
    assert p.calls[0][1] == [0] * 10                   # ten children share one series (:60-66, :88)
    assert (r.get_avg(), r.get_std()) == (10.0, 0.0)
