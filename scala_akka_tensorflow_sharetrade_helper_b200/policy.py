"""Host-side mirror of the reference's decision-policy actor over the C ABI.

Same vocabulary as src/main/scala/QDecisionPolicyActor.scala: the messages SelectionAction / UpdateQ /
Updated (:31-34), the actions Buy / Sell / Hold (:26-29, index order :17) and the size guards that
throw IllegalArgumentException (:56-57, :64-65).  Where the reference sends one message per
episode through the actor's mailbox, this mirror accepts a batch of rows and makes one GPU launch.
All arithmetic happens in libqpolicy.so; nothing here computes.
"""
from __future__ import annotations

import ctypes as C
import enum
from dataclasses import dataclass

import numpy as np

from . import _capi


class IllegalArgumentException(ValueError):
    """QP_EINVAL: what the reference throws inside `receive` on a size mismatch."""


class QPolicyError(RuntimeError):
    pass


class Action(enum.IntEnum):
    """QDecisionPolicyActor.scala:17 — `val actions = Seq(Buy, Sell, Hold)`."""
    Buy = 0
    Sell = 1
    Hold = 2


Buy, Sell, Hold = Action.Buy, Action.Sell, Action.Hold


@dataclass
class SelectionAction:
    """QDecisionPolicyActor.scala:32."""
    currentState: np.ndarray
    step: float


@dataclass
class UpdateQ:
    """QDecisionPolicyActor.scala:33."""
    state: np.ndarray
    reward: float
    nextState: np.ndarray


class Updated:
    """QDecisionPolicyActor.scala:34."""


def check(status: int) -> None:
    if status == _capi.QP_OK:
        return
    msg = _capi.last_error()
    if status == _capi.QP_EINVAL:
        raise IllegalArgumentException(msg)
    raise QPolicyError(f"qpolicy status {status}: {msg}")


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


@dataclass
class PolicyConfig:
    """The reference's compile-time constants (QDecisionPolicyActor.scala:17-22, :50)."""
    window: int = 201
    hidden: int = 200
    epsilon: float = 0.9
    gamma: float = 0.001
    lr: float = 0.01
    bias: float = 0.1
    step_div: float = 1000.0
    adagrad_init_acc: float = 1e-8
    bookkeeping: int = 0   # 0 = REFERENCE, 1 = RLPY
    device: int = 0
    seed: int = 42

    def c(self) -> _capi.QpConfig:
        return _capi.QpConfig(self.window, self.hidden, self.epsilon, self.gamma, self.lr, self.bias,
                              self.step_div, self.adagrad_init_acc, self.bookkeeping, self.device, self.seed)


class QDecisionPolicy:
    """The policy actor's body: one Q-network, its AdaGrad state, and (optionally) episodes."""

    def __init__(self, config: PolicyConfig | None = None):
        self.config = config or PolicyConfig()
        self._lib = _capi.lib()
        self._h = C.c_void_p()
        cfg = self.config.c()
        check(self._lib.qp_create(C.byref(cfg), C.byref(self._h)))
        self.input_dim = int(self._lib.qp_input_dim(self._h))
        self.param_count = int(self._lib.qp_param_count(self._h))
        self.iteration = 0  # QDecisionPolicyActor.scala:76

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.qp_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- weights (no reference counterpart; needed for parity and checkpoints) -----------------
    def set_weights(self, params=None, acc=None):
        p = None if params is None else _f32(params).reshape(-1)
        a = None if acc is None else _f32(acc).reshape(-1)
        for v in (p, a):
            if v is not None and v.size != self.param_count:
                raise IllegalArgumentException(f"expected {self.param_count} parameters, got {v.size}")
        check(self._lib.qp_set_weights(self._h, _ptr(p), _ptr(a)))

    def get_weights(self):
        p = np.empty(self.param_count, np.float32)
        a = np.empty(self.param_count, np.float32)
        check(self._lib.qp_get_weights(self._h, _ptr(p), _ptr(a)))
        return p, a

    def last_grad(self):
        g = np.empty(self.param_count, np.float32)
        check(self._lib.qp_get_last_grad(self._h, _ptr(g)))
        return g

    # ---- message level ---------------------------------------------------------------------------
    def select_action(self, states, steps, env_ids=None, return_q=False):
        """SelectionAction for a batch of rows; returns int32 action indices (and q if asked)."""
        states = _f32(states)
        if states.ndim == 1:
            states = states[None, :]
        B, L = states.shape
        steps = np.ascontiguousarray(np.broadcast_to(np.asarray(steps, np.float32), (B,)))
        ids = None if env_ids is None else np.ascontiguousarray(env_ids, np.int64)
        actions = np.empty(B, np.int32)
        q = np.empty((B, 3), np.float32)
        check(self._lib.qp_select_action(self._h, _ptr(states), L, _ptr(steps), _ptr(ids), _ptr(actions), _ptr(q), B))
        return (actions, q) if return_q else actions

    def update_q(self, states, rewards, next_states):
        """UpdateQ for a batch of rows: gradient SUM, one AdaGrad step.  Returns Updated."""
        states = _f32(states)
        next_states = _f32(next_states)
        if states.ndim == 1:
            states = states[None, :]
        if next_states.ndim == 1:
            next_states = next_states[None, :]
        B = states.shape[0]
        if next_states.shape[0] != B:
            raise IllegalArgumentException("state and nextState batch sizes differ")
        rewards = np.ascontiguousarray(np.broadcast_to(np.asarray(rewards, np.float32), (B,)))
        check(self._lib.qp_update_q(self._h, _ptr(states), states.shape[1], _ptr(rewards),
                                    _ptr(next_states), next_states.shape[1], B))
        self.iteration += 1
        return Updated()

    def receive(self, msg):
        """Message-style entry, as the actor's `receive` (QDecisionPolicyActor.scala:54-77)."""
        if isinstance(msg, SelectionAction):
            return Action(int(self.select_action(np.asarray(msg.currentState).reshape(1, -1), msg.step)[0]))
        if isinstance(msg, UpdateQ):
            return self.update_q(np.asarray(msg.state).reshape(1, -1), msg.reward,
                                 np.asarray(msg.nextState).reshape(1, -1))
        raise TypeError(f"unhandled message {msg!r}")

    # ---- episode level ---------------------------------------------------------------------------
    def env_create(self, prices, series_of_env=None, budget=1000.0, shares=0, env_id_base=0):
        prices = _f32(prices)
        if prices.ndim == 1:
            prices = prices[None, :]
        n_series, T = prices.shape
        if series_of_env is None:
            series_of_env = np.arange(n_series)
        soe = np.ascontiguousarray(series_of_env, np.int32)
        check(self._lib.qp_env_create(self._h, _ptr(prices), T, n_series, _ptr(soe), soe.size,
                                      int(env_id_base), float(budget), int(shares)))
        self.n_envs = int(soe.size)

    def env_reset(self):
        check(self._lib.qp_env_reset(self._h))

    @property
    def steps_total(self):
        return int(self._lib.qp_env_steps_total(self._h))

    @property
    def steps_done(self):
        return int(self._lib.qp_env_steps_done(self._h))

    def env_step(self, n_steps=1, sync=True):
        fn = self._lib.qp_env_step if sync else self._lib.qp_env_step_async
        check(fn(self._h, int(n_steps)))

    def sync(self):
        check(self._lib.qp_sync(self._h))

    def set_option(self, key: str, value: int):
        check(self._lib.qp_set_option(self._h, key.encode(), int(value)))

    def get_option(self, key: str) -> int:
        v = C.c_int64()
        check(self._lib.qp_get_option(self._h, key.encode(), C.byref(v)))
        return int(v.value)

    def phase_clocks(self, n_steps):
        out = np.zeros((n_steps, 16), np.int64)
        check(self._lib.qp_debug_phase_clocks(self._h, int(n_steps), _ptr(out)))
        return out

    def env_step_timed(self, n_steps):
        ms = np.zeros(_capi.QP_N_KERNELS, np.float32)
        check(self._lib.qp_env_step_timed(self._h, int(n_steps), _ptr(ms)))
        return ms

    def env_step_trace(self, n_steps):
        n = self.n_envs
        rec = dict(
            action_sel=np.empty((n_steps, n), np.int32), action_eff=np.empty((n_steps, n), np.int32),
            explored=np.empty((n_steps, n), np.int32), k_next=np.empty((n_steps, n), np.int32),
            reward=np.empty((n_steps, n), np.float32), q_s=np.empty((n_steps, n, 3), np.float32),
            q_n=np.empty((n_steps, n, 3), np.float32), target=np.empty((n_steps, n), np.float32),
            budget=np.empty((n_steps, n), np.float64), shares=np.empty((n_steps, n), np.int32),
        )
        t = _capi.QpTrace(*[_ptr(rec[k]) for k in ("action_sel", "action_eff", "explored", "k_next", "reward",
                                                   "q_s", "q_n", "target", "budget", "shares")])
        check(self._lib.qp_env_step_trace(self._h, int(n_steps), C.byref(t)))
        return rec

    def portfolios(self):
        out = np.empty(self.n_envs, np.float64)
        check(self._lib.qp_env_get_portfolio(self._h, _ptr(out)))
        return out

    def env_state(self):
        b = np.empty(self.n_envs, np.float64)
        s = np.empty(self.n_envs, np.int32)
        v = np.empty(self.n_envs, np.float64)
        check(self._lib.qp_env_get_state(self._h, _ptr(b), _ptr(s), _ptr(v)))
        return b, s, v

    def env_set_state(self, budget=None, shares=None, share_value=None, step=0):
        b = None if budget is None else np.ascontiguousarray(budget, np.float64)
        s = None if shares is None else np.ascontiguousarray(shares, np.int32)
        v = None if share_value is None else np.ascontiguousarray(share_value, np.float64)
        check(self._lib.qp_env_set_state(self._h, _ptr(b), _ptr(s), _ptr(v), int(step)))

    def portfolio_stats(self):
        m, s = C.c_double(), C.c_double()
        check(self._lib.qp_env_portfolio_stats(self._h, C.byref(m), C.byref(s)))
        return m.value, s.value

    def train(self, prices, series_of_env=None, budget=1000.0, shares=0, env_id_base=0):
        """TrainerChildActor.train (:64-71) for every episode, host buffers in, portfolios out."""
        prices = _f32(prices)
        if prices.ndim == 1:
            prices = prices[None, :]
        n_series, T = prices.shape
        if series_of_env is None:
            series_of_env = np.arange(n_series)
        soe = np.ascontiguousarray(series_of_env, np.int32)
        out = np.empty(soe.size, np.float64)
        check(self._lib.qp_train(self._h, _ptr(prices), T, n_series, _ptr(soe), soe.size, int(env_id_base),
                                 float(budget), int(shares), _ptr(out)))
        self.n_envs = int(soe.size)
        return out

    def dist_init(self, dist, rank: int, world: int, p2p: bool = True, nvls: bool = True):
        """Join the gradient all-reduce group.  `dist` is an initialised torch.distributed (any backend):
        it only carries the 128-byte NCCL rendezvous id from rank 0 to the others, and the handles of the
        exchange blocks.  p2p: the all-reduce runs inside the persistent kernel over NVLink peer memory.
        nvls: the exchange blocks are torch symmetric memory with a multicast (NVLS) mapping where the
        system offers one, so that a share is ONE store the NVSwitch delivers to every rank; otherwise
        (or with nvls=False) plain cudaMalloc blocks shared by CUDA IPC and world - 1 stores per share."""
        import torch
        buf = (C.c_ubyte * 128)()
        if rank == 0:
            check(self._lib.qp_dist_unique_id(buf, 128))
        ids = [bytes(buf)]
        if world > 1:
            dist.broadcast_object_list(ids, src=0)
        uid = (C.c_ubyte * 128).from_buffer_copy(ids[0])
        check(self._lib.qp_dist_init(self._h, uid, int(rank), int(world)))
        self.p2p = False
        self.nvls = False
        self._symm = None
        if world > 1 and p2p:
            ok = False
            if nvls:
                ok = self._attach_symmetric_memory(torch, dist, world)
                flags = [None] * world
                dist.all_gather_object(flags, bool(ok))
                ok = all(flags)
                if not ok and self._symm is not None:
                    raise QPolicyError("symmetric-memory exchange blocks attached on some ranks only")
                self.nvls = ok and self._symm[2] != 0
            if not ok:
                # exchange CUDA IPC handles of cudaMalloc'd blocks
                hbuf = (C.c_ubyte * 64)()
                ok = self._lib.qp_dist_p2p_export(self._h, hbuf, 64) == _capi.QP_OK
                gathered = [None] * world
                dist.all_gather_object(gathered, (ok, bytes(hbuf)))
                if all(g[0] for g in gathered):
                    blob = b"".join(g[1] for g in gathered)
                    ok = self._lib.qp_dist_p2p_import(self._h, (C.c_ubyte * len(blob)).from_buffer_copy(blob), world) == _capi.QP_OK
                else:
                    ok = False
                flags = [None] * world
                dist.all_gather_object(flags, bool(ok))
                ok = all(flags)
            self.p2p = ok
            self.set_option("p2p", 1 if self.p2p else 0)   # every rank must take the same path

    def _attach_symmetric_memory(self, torch, dist, world) -> bool:
        """Exchange blocks as torch symmetric memory (peer-mapped by the runtime, with a multicast address where
        NVLS is available).  Collective; returns False (and leaves nothing attached) when this torch / system
        cannot do it."""
        try:
            import torch.distributed._symmetric_memory as symm_mem
            nbytes = int(self._lib.qp_dist_p2p_bytes(self._h))
            dev = torch.device("cuda", self.config.device)
            t = symm_mem.empty(nbytes, dtype=torch.uint8, device=dev)
            t.zero_()
            torch.cuda.synchronize(dev)
            hdl = symm_mem.rendezvous(t, dist.group.WORLD)
            ptrs = (C.c_uint64 * world)(*[int(p_) for p_ in hdl.buffer_ptrs])
            mc = int(hdl.multicast_ptr) if getattr(hdl, "has_multicast_support", False) or hdl.multicast_ptr else 0
            if self._lib.qp_dist_p2p_attach(self._h, ptrs, world, C.c_uint64(mc)) != _capi.QP_OK:
                return False
            self._symm = (t, hdl, mc)          # kept alive as long as the policy
            return True
        except Exception:
            return False

    # ---- checkpoint (the reference's saveSnapshot stub, QDecisionPolicyActor.scala:74, :91-93) ------
    def save_checkpoint(self, path: str):
        check(self._lib.qp_save_checkpoint(self._h, str(path).encode()))

    def load_checkpoint(self, path: str):
        check(self._lib.qp_load_checkpoint(self._h, str(path).encode()))
        self.iteration = int(self._lib.qp_iteration(self._h))

    def set_snapshot(self, path, every: int = 500):
        check(self._lib.qp_set_snapshot(self._h, None if path is None else str(path).encode(), int(every)))

    def measure_fp32_peak(self) -> float:
        t = C.c_double()
        check(self._lib.qp_measure_fp32_peak(self._h, C.byref(t)))
        return t.value

    def measure_fp32_outer_peak(self) -> float:
        t = C.c_double()
        check(self._lib.qp_measure_fp32_outer_peak(self._h, C.byref(t)))
        return t.value

    @property
    def stream(self) -> int:
        return int(self._lib.qp_stream(self._h) or 0)

    @property
    def launch_count(self) -> int:
        return int(self._lib.qp_launch_count(self._h))


class BatchingPolicy:
    """Front-end for many child actors sharing one policy (TrainerRouterActor.scala:44, :60-66): each
    child enqueues its SelectionAction / UpdateQ; flush() answers all of them with one GPU launch per
    kind.  enqueue_* are thread-safe and do not touch the GPU."""

    def __init__(self, policy: QDecisionPolicy):
        self.policy = policy
        self._lib = policy._lib
        self._h = policy._h

    def enqueue_select(self, state, step: float, env_id: int) -> int:
        state = _f32(state).reshape(-1)
        t = C.c_int64()
        check(self._lib.qp_enqueue_select(self._h, _ptr(state), state.size, float(step), int(env_id), C.byref(t)))
        return int(t.value)

    def enqueue_update(self, state, reward: float, next_state) -> None:
        state = _f32(state).reshape(-1)
        next_state = _f32(next_state).reshape(-1)
        check(self._lib.qp_enqueue_update(self._h, _ptr(state), state.size, float(reward), _ptr(next_state), next_state.size))

    def flush(self):
        ns, nu = C.c_int64(), C.c_int64()
        check(self._lib.qp_flush(self._h, C.byref(ns), C.byref(nu)))
        if nu.value:
            self.policy.iteration += 1
        return int(ns.value), int(nu.value)

    def action(self, ticket: int) -> Action:
        a = C.c_int32()
        check(self._lib.qp_take_action(self._h, int(ticket), C.byref(a)))
        return Action(int(a.value))

    @property
    def pending(self) -> int:
        return int(self._lib.qp_pending(self._h))
