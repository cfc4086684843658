"""ctypes front-end of the CPU oracle (oracle/qp_oracle.c).

TEST INFRASTRUCTURE ONLY — see the header of qp_oracle.c.  Imported by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs, never by the
product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ACTIONS = 3
BUY, SELL, HOLD = 0, 1, 2
REFERENCE, RLPY = 0, 1


class _Cfg(C.Structure):
    _fields_ = [
        ("wn", C.c_int32), ("hidden", C.c_int32), ("epsilon", C.c_float), ("gamma", C.c_float),
        ("lr", C.c_float), ("bias", C.c_float), ("step_div", C.c_float),
        ("bookkeeping", C.c_int32), ("seed", C.c_uint64),
    ]


class _Trace(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("action_sel", "action_eff", "explored", "k_next", "reward", "q_s", "q_n", "target")]


class _Force(C.Structure):
    _fields_ = [("action", C.c_void_p), ("k_next", C.c_void_p)]


class _EnvState(C.Structure):
    _fields_ = [("budget", C.c_void_p), ("shares", C.c_void_p), ("share_value", C.c_void_p)]


def build(force: bool = False) -> None:
    """Compile the oracle's two shared objects (fp32 oracle, fp64 twin) with the committed Makefile."""
    need = force or not all(os.path.exists(os.path.join(_HERE, f))
                            for f in ("libqp_oracle_f32.so", "libqp_oracle_f64.so"))
    src = os.path.join(_HERE, "qp_oracle.c")
    if not need and os.path.exists(src):
        need = any(os.path.getmtime(os.path.join(_HERE, f)) < os.path.getmtime(src)
                   for f in ("libqp_oracle_f32.so", "libqp_oracle_f64.so"))
    if need:
        subprocess.run(["make", "-C", _HERE, "-s", "-B"], check=True)


_libs: dict = {}


def _lib(dtype):
    key = np.dtype(dtype).name
    if key not in _libs:
        build()
        name = {"float32": "libqp_oracle_f32.so", "float64": "libqp_oracle_f64.so"}[key]
        lib = C.CDLL(os.path.join(_HERE, name))
        assert lib.qpo_sizeof_real() == np.dtype(dtype).itemsize
        lib.qpo_select_action.restype = C.c_int
        lib.qpo_max_threads.restype = C.c_int
        _libs[key] = lib
    return _libs[key]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


@dataclass
class Config:
    wn: int = 201            # QDecisionPolicyActor.scala:18 (201 prices + budget + shares)
    hidden: int = 200        # :22
    epsilon: float = 0.9     # :19
    gamma: float = 0.001     # :20
    lr: float = 0.01         # :50
    bias: float = 0.1        # :42, :46
    step_div: float = 1000.0  # :58
    bookkeeping: int = REFERENCE
    seed: int = 42
    adagrad_init_acc: float = 1e-8

    @property
    def D(self):
        return self.wn + 2

    @property
    def P(self):
        return self.D * self.hidden + self.hidden * ACTIONS

    def c(self):
        return _Cfg(self.wn, self.hidden, self.epsilon, self.gamma, self.lr, self.bias,
                    self.step_div, self.bookkeeping, self.seed)


def philox4x32_10(ctr, key):
    lib = _lib(np.float32)
    c = (C.c_uint32 * 4)(*ctr)
    k = (C.c_uint32 * 2)(*key)
    o = (C.c_uint32 * 4)()
    lib.qpo_philox4x32_10(c, k, o)
    return [int(v) for v in o]


def gen_prices(seed: int, first_series: int, n_series: int, T: int) -> np.ndarray:
    out = np.empty((n_series, T), np.float32)
    _lib(np.float32).qpo_gen_prices(C.c_uint64(seed), C.c_int64(first_series), C.c_int64(n_series),
                                    C.c_int32(T), _p(out))
    return out


def gen_normal(seed: int, first: int, n: int) -> np.ndarray:
    out = np.empty(n, np.float32)
    _lib(np.float32).qpo_gen_normal(C.c_uint64(seed), C.c_int64(first), C.c_int64(n), _p(out))
    return out


def portfolio_stats(x):
    x = np.ascontiguousarray(x, np.float64)
    m, s = C.c_double(), C.c_double()
    _lib(np.float32).qpo_portfolio_stats(_p(x), C.c_int64(x.size), C.byref(m), C.byref(s))
    return m.value, s.value


def max_threads() -> int:
    return _lib(np.float32).qpo_max_threads()


class Oracle:
    """One policy (weights + AdaGrad accumulators) and, optionally, a set of episodes."""

    def __init__(self, cfg: Config, dtype=np.float32, params=None, acc=None):
        self.cfg = cfg
        self.dtype = np.dtype(dtype)
        self.lib = _lib(dtype)
        if params is None:
            params = gen_normal(cfg.seed, 0, cfg.P)
        self.params = np.ascontiguousarray(params, self.dtype).copy()
        if acc is None:
            acc = np.full(cfg.P, cfg.adagrad_init_acc, np.float32)
        self.acc = np.ascontiguousarray(acc, self.dtype).copy()
        self.env = None

    # -- message level -----------------------------------------------------------------------
    def forward(self, x):
        x = np.ascontiguousarray(x, self.dtype).reshape(-1, self.cfg.D)
        h = np.empty((x.shape[0], self.cfg.hidden), self.dtype)
        q = np.empty((x.shape[0], ACTIONS), self.dtype)
        cc = self.cfg.c()
        for b in range(x.shape[0]):
            self.lib.qpo_forward(C.byref(cc), _p(self.params), _p(x[b]), _p(h[b]), _p(q[b]))
        return h, q

    def select_action(self, x, step, env_ids=None):
        x = np.ascontiguousarray(x, self.dtype).reshape(-1, self.cfg.D)
        B = x.shape[0]
        step = np.broadcast_to(np.asarray(step, np.float32), (B,))
        env_ids = np.arange(B) if env_ids is None else np.asarray(env_ids)
        q = np.empty((B, ACTIONS), self.dtype)
        act = np.empty(B, np.int32)
        exp = np.empty(B, np.int32)
        cc = self.cfg.c()
        e = C.c_int()
        for b in range(B):
            act[b] = self.lib.qpo_select_action(C.byref(cc), _p(self.params), _p(x[b]),
                                                C.c_float(float(step[b])), C.c_uint64(int(env_ids[b])),
                                                _p(q[b]), C.byref(e))
            exp[b] = e.value
        return act, q, exp

    def td_grad(self, s, reward, s2):
        s = np.ascontiguousarray(s, self.dtype).reshape(-1, self.cfg.D)
        s2 = np.ascontiguousarray(s2, self.dtype).reshape(-1, self.cfg.D)
        B = s.shape[0]
        reward = np.ascontiguousarray(np.broadcast_to(np.asarray(reward, np.float32), (B,)))
        g = np.empty(self.cfg.P, self.dtype)
        q_s = np.empty((B, ACTIONS), self.dtype)
        q_n = np.empty((B, ACTIONS), self.dtype)
        k = np.empty(B, np.int32)
        tgt = np.empty(B, self.dtype)
        cc = self.cfg.c()
        self.lib.qpo_td_grad_batch(C.byref(cc), _p(self.params), _p(s), _p(reward), _p(s2),
                                   C.c_int64(B), _p(g), _p(q_s), _p(q_n), _p(k), _p(tgt))
        return g, dict(q_s=q_s, q_n=q_n, k_next=k, target=tgt)

    def update_q(self, s, reward, s2):
        s = np.ascontiguousarray(s, self.dtype).reshape(-1, self.cfg.D)
        s2 = np.ascontiguousarray(s2, self.dtype).reshape(-1, self.cfg.D)
        B = s.shape[0]
        reward = np.ascontiguousarray(np.broadcast_to(np.asarray(reward, np.float32), (B,)))
        g = np.empty(self.cfg.P, self.dtype)
        cc = self.cfg.c()
        self.lib.qpo_update_q(C.byref(cc), _p(self.params), _p(self.acc), _p(s), _p(reward), _p(s2),
                              C.c_int64(B), _p(g))
        return g

    # -- episode level -----------------------------------------------------------------------
    def env_create(self, prices, series_of_env, budget0: float, shares0: int, env_id_base: int = 0):
        prices = np.ascontiguousarray(prices, np.float32)
        if prices.ndim == 1:
            prices = prices[None, :]
        if prices.shape[1] <= self.cfg.wn:
            # TrainerChildActor.scala:65, :70
            raise ValueError("Stock price count should be more than Tensorflow input nodes")
        soe = np.ascontiguousarray(series_of_env, np.int32)
        n = soe.size
        self.env = dict(
            prices=prices, soe=soe, n=n, B0=float(budget0), N0=int(shares0), base=int(env_id_base),
            budget=np.full(n, float(budget0), np.float64), shares=np.full(n, int(shares0), np.int32),
            share_value=np.zeros(n, np.float64), step=0,
        )

    @property
    def n_steps_total(self):
        return self.env["prices"].shape[1] - self.cfg.wn  # TrainerChildActor.scala:67

    def env_step(self, n_steps=1, trace=False, mode=0, want_grad=False, threads=0, force_action=None,
                 force_knext=None):
        """force_action / force_knext: int32 per episode, entries >= 0 replace the oracle's own choice of
        the selected action / of argmax q(s') for that episode (teacher hook of the parity tests, used only
        where the unforced run showed the differing choice to lie inside the stated tolerance)."""
        ev = self.env
        fc = None
        if force_action is not None or force_knext is not None:
            assert n_steps == 1 and not threads
            fa = None if force_action is None else np.ascontiguousarray(force_action, np.int32)
            fk = None if force_knext is None else np.ascontiguousarray(force_knext, np.int32)
            fc = _Force(_p(fa), _p(fk))
        n = ev["n"]
        traces = []
        g = np.empty(self.cfg.P, self.dtype) if want_grad else None
        partials = np.empty((max(threads, 1), self.cfg.P), self.dtype) if threads else None
        cc = self.cfg.c()
        st = _EnvState(_p(ev["budget"]), _p(ev["shares"]), _p(ev["share_value"]))
        for _ in range(n_steps):
            i = ev["step"]
            if i >= self.n_steps_total:
                raise ValueError("episode finished")
            tr = None
            rec = None
            if trace:
                rec = dict(
                    action_sel=np.empty(n, np.int32), action_eff=np.empty(n, np.int32),
                    explored=np.empty(n, np.int32), k_next=np.empty(n, np.int32),
                    reward=np.empty(n, np.float32), q_s=np.empty((n, ACTIONS), self.dtype),
                    q_n=np.empty((n, ACTIONS), self.dtype), target=np.empty(n, self.dtype),
                )
                tr = _Trace(*[_p(rec[k]) for k in ("action_sel", "action_eff", "explored", "k_next",
                                                   "reward", "q_s", "q_n", "target")])
            if threads:
                self.lib.qpo_env_step_mt(C.byref(cc), _p(self.params), _p(self.acc), _p(ev["prices"]),
                                         C.c_int32(ev["prices"].shape[1]), _p(ev["soe"]), C.c_int64(n),
                                         C.c_int64(ev["base"]), C.c_int32(i), C.c_double(ev["B0"]),
                                         C.c_int32(ev["N0"]), C.byref(st), C.c_int(threads), _p(partials))
            else:
                self.lib.qpo_env_step_forced(C.byref(cc), _p(self.params), _p(self.acc), _p(ev["prices"]),
                                             C.c_int32(ev["prices"].shape[1]), _p(ev["soe"]), C.c_int64(n),
                                             C.c_int64(ev["base"]), C.c_int32(i), C.c_double(ev["B0"]),
                                             C.c_int32(ev["N0"]), C.byref(st),
                                             C.byref(tr) if tr is not None else None, C.c_int(mode), _p(g),
                                             C.byref(fc) if fc is not None else None)
            ev["step"] = i + 1
            if rec is not None:
                rec["budget"] = ev["budget"].copy()
                rec["shares"] = ev["shares"].copy()
                rec["share_value"] = ev["share_value"].copy()
                if g is not None:
                    rec["grad"] = g.copy()
                traces.append(rec)
        return traces

    def adagrad(self, g):
        """Apply one ResourceApplyAdagrad step with an externally summed gradient (multi-rank tests)."""
        g = np.ascontiguousarray(g, self.dtype)
        cc = self.cfg.c()
        self.lib.qpo_adagrad(C.byref(cc), _p(self.params), _p(self.acc), _p(g), C.c_int64(self.cfg.P))

    def portfolios(self):
        ev = self.env
        return ev["budget"] + ev["shares"] * ev["share_value"]  # TrainerChildActor.scala:68
