"""Host-side mirror of the reference's episode actor and router over the C ABI.

TrainerChild keeps the FSM contract of src/main/scala/TrainerChildActor.scala (:30-62): states
Ready / Trained, messages Train / GetPortfolio / Initialise, replies Trained / TrainedData /
NotComputed / Initialised.  Its train() body — the fold at :82-103 — runs on the GPU.  TrainerRouter
mirrors GetAvg / GetStd of TrainerRouterActor.scala (:89-94, :149-151) over many episodes that share
one policy (:44, :60-66).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .policy import IllegalArgumentException, QDecisionPolicy


class NotComputed:
    """TrainerRouterActor.NotComputed: no portfolio yet."""
    def __eq__(self, other):
        return isinstance(other, NotComputed)


class Trained:
    def __eq__(self, other):
        return isinstance(other, Trained)


class Initialised:
    def __eq__(self, other):
        return isinstance(other, Initialised)


@dataclass
class TrainedData:
    portfolio: float


@dataclass
class StockDataResponse:
    """SharePriceGetter.StockDataResponse(name, TreeMap[LocalDate, Double]) (SharePriceGetter.scala:14-15).
    `prices` maps an orderable date key to a price; iteration is by ascending key, like a TreeMap."""
    name: str
    prices: dict

    def sorted_prices(self) -> np.ndarray:
        return np.asarray([self.prices[k] for k in sorted(self.prices)], dtype=np.float64).astype(np.float32)


class TrainerChild:
    """One episode actor: TrainerChildActor(policyActor, myBudget, noOfStocks)."""

    def __init__(self, policy: QDecisionPolicy, my_budget: float, no_of_stocks: int, env_id: int = 0):
        self.policy = policy
        self.my_budget = float(my_budget)
        self.no_of_stocks = int(no_of_stocks)
        self.env_id = int(env_id)
        self.state = "Ready"
        self.data = NotComputed()

    def train(self, stock_data: StockDataResponse) -> float:
        """TrainerChildActor.train (:64-71).  Too few prices -> IllegalArgumentException (:70)."""
        prices = stock_data.sorted_prices()
        if not prices.size > self.policy.config.window:
            raise IllegalArgumentException("Stock price count should be more than Tensorflow input nodes")
        out = self.policy.train(prices[None, :], [0], self.my_budget, self.no_of_stocks, self.env_id)
        return float(out[0])

    # -- FSM (:32-60) --------------------------------------------------------------------------
    def receive(self, msg, stock_data: StockDataResponse | None = None):
        if self.state == "Ready":
            if msg == "Train":
                self.data = TrainedData(self.train(stock_data))
                self.state = "Trained"
                return Trained()
            if msg == "GetPortfolio":
                return NotComputed()
            if msg == "Initialise":
                return None  # "already in ready state"
        else:
            if msg == "Train":
                return None  # "it's already trained"
            if msg == "GetPortfolio":
                return self.data
            if msg == "Initialise":
                self.state = "Ready"
                self.data = NotComputed()
                return Initialised()
        raise TypeError(f"unhandled message {msg!r}")


class TrainerRouter:
    """Fan-out of many episodes over one shared policy, batched into one GPU job.

    The reference broadcasts Train to 10 children that interleave through the policy's mailbox
    (TrainerRouterActor.scala:36, :60-66, :88); here the children's steps run in lock-step, the
    gradient is summed over them, and GetAvg / GetStd are reduced on the device.
    """

    def __init__(self, policy: QDecisionPolicy, n_children: int = 10, budget: float = 2400.0, shares: int = 0):
        self.policy = policy
        self.n_children = int(n_children)
        self.budget = float(budget)
        self.shares = int(shares)
        self.stock_data = None
        self.trained = False

    def send_training_data(self, stock_data: StockDataResponse):
        self.stock_data = stock_data

    def start_training(self):
        if self.stock_data is None:
            return "NoTrainingDataReceived"
        prices = self.stock_data.sorted_prices()
        if not prices.size > self.policy.config.window:
            raise IllegalArgumentException("Stock price count should be more than Tensorflow input nodes")
        self.policy.train(prices[None, :], np.zeros(self.n_children, np.int32), self.budget, self.shares)
        self.trained = True
        return "Completed"

    def is_everything_done(self):
        return "Completed" if self.trained else "TrainingNotCompleted"

    def get_avg(self):
        return self.policy.portfolio_stats()[0] if self.trained else NotComputed()

    def get_std(self):
        return self.policy.portfolio_stats()[1] if self.trained else NotComputed()
