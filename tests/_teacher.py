"""Teacher-forced comparison of ONE env-step of every episode: device (through the C ABI) against the oracle.

Both sides start the step from the oracle's weights, accumulators and episode state.  What the step
produces is compared under the stated tolerance (tests/_tol.py): Q(s), Q(s') and the TD target within
tolerance; explore flags, effective actions, share counts, budgets and rewards bit for bit.

An exploit action (argmax Q(s)) or the TD index (argmax Q(s')) may differ from the oracle's only where
the oracle's top-two gap is inside the Q tolerance.  Where that happens the oracle is run a second
time from the same start with the device's choice imposed on those episodes (oracle.env_step's
force_action / force_knext), so that book-keeping, gradient, weights and carry are compared in EVERY
case instead of being skipped; the number of imposed choices is returned and bounded by the callers.
"""
import numpy as np

from oracle import oracle as O
from tests import _tol


def states_of(cfg, prices, soe, i, budget, shares):
    w = np.stack([prices[s, i:i + cfg.wn] for s in soe]) if len(soe) < 4096 else prices[np.asarray(soe), i:i + cfg.wn]
    return np.concatenate([w, np.asarray(budget, np.float64).astype(np.float32)[:, None],
                           np.asarray(shares, np.float32)[:, None]], axis=1)


class Tally:
    """How many teacher-forced steps were compared in full, and how many choices had to be imposed."""

    def __init__(self):
        self.steps = 0
        self.rows = 0
        self.forced_actions = 0
        self.forced_knext = 0
        self.grad_checked = 0
        self.kink_columns = 0
        self.grad_err_vs_twin = []      # (device, fp32 oracle) relative Frobenius distance from the fp64 twin

    def assert_mostly_unforced(self, max_frac=0.02):
        assert self.steps > 0 and self.grad_checked == self.steps, "a teacher-forced step was not compared in full"
        frac = (self.forced_actions + self.forced_knext) / max(1, self.rows)
        assert frac <= max_frac, (f"{self.forced_actions} actions and {self.forced_knext} TD indices of {self.rows} rows "
                                  f"had to be imposed on the oracle: more than {max_frac:.0%}")


def _snapshot(o):
    ev = o.env
    return (o.params.copy(), o.acc.copy(), ev["budget"].copy(), ev["shares"].copy(), ev["share_value"].copy(), ev["step"])


def _restore(o, snap):
    o.params[:], o.acc[:] = snap[0], snap[1]
    o.env["budget"][:], o.env["shares"][:], o.env["share_value"][:] = snap[2], snap[3], snap[4]
    o.env["step"] = snap[5]


def teacher_forced_step(cfg, o, pol, prices, soe, i, what, tally=None, n_terms=1, check_state=True):
    """One step at env-step index i.  `o` (oracle.Oracle with episodes) supplies the start state; on return
    it stands after the step (with the device's tolerated choices imposed, if any)."""
    o.env["step"] = i
    snap = _snapshot(o)
    w, acc, b, s, v, _ = snap
    pol.set_weights(w, acc)
    pol.env_set_state(b, s, v, step=i)
    x = states_of(cfg, prices, soe, i, b, s)
    ref = o.env_step(1, trace=True, want_grad=True)[0]
    tr = {k: a[0] for k, a in pol.env_step_trace(1).items()}

    # ---- SelectionAction (QDecisionPolicyActor.scala:56-62)
    _tol.assert_q_close(cfg, w, x, tr["q_s"], ref["q_s"], what + " q_s")
    np.testing.assert_array_equal(tr["explored"], ref["explored"], err_msg=what + " explored")
    n_act = _tol.assert_actions_match(cfg, w, x, tr["action_sel"], ref["action_sel"], ref["q_s"], what + " action")
    force_a = None
    if n_act:
        force_a = np.where(tr["action_sel"] != ref["action_sel"], tr["action_sel"], -1).astype(np.int32)
        _restore(o, snap)
        ref = o.env_step(1, trace=True, want_grad=True, force_action=force_a)[0]
        np.testing.assert_array_equal(tr["action_sel"], ref["action_sel"])

    # ---- decision, reward, carry (TrainerChildActor.scala:92-102, 118-123, 136-146): bit-exact on every row
    for k in ("action_eff", "shares", "budget", "reward"):
        np.testing.assert_array_equal(tr[k], ref[k], err_msg=f"{what} {k}")

    # ---- UpdateQ (QDecisionPolicyActor.scala:64-73): Q(s'), its argmax, the target
    x2 = states_of(cfg, prices, soe, i + 1, ref["budget"], ref["shares"])
    _tol.assert_q_close(cfg, w, x2, tr["q_n"], ref["q_n"], what + " q_n")
    n_k = _tol.assert_actions_match(cfg, w, x2, tr["k_next"], ref["k_next"], ref["q_n"], what + " k_next")
    if n_k:
        force_k = np.where(tr["k_next"] != ref["k_next"], tr["k_next"], -1).astype(np.int32)
        _restore(o, snap)
        ref = o.env_step(1, trace=True, want_grad=True, force_action=force_a, force_knext=force_k)[0]
        np.testing.assert_array_equal(tr["k_next"], ref["k_next"])
    tol_t = abs(cfg.gamma) * _tol.q_tol(cfg, w, x2, ref["q_n"]).max(axis=-1) + 1e-6 * np.maximum(1.0, np.abs(ref["target"]))
    assert (np.abs(tr["target"].astype(np.float64) - ref["target"]) <= tol_t).all(), what + " target"

    # ---- gradient sum, AdaGrad (QDecisionPolicyActor.scala:49-50, :73), carry
    g = pol.last_grad()
    # The gradient is discontinuous where a hidden unit sits on its ReLU kink (ReluGrad passes where the output is
    # strictly positive): a unit whose pre-activation is within fp32 rounding of zero for some episode may be masked
    # differently by two correct implementations, which moves that column of dW1 by a whole episode's contribution.
    # Such columns (rare; counted) are left out of the gradient comparison.
    D, H = cfg.D, cfg.hidden
    x64 = x.astype(np.float64)
    W1 = w[:D * H].astype(np.float64).reshape(D, H)
    z = x64 @ W1 + cfg.bias
    zc = np.abs(x64) @ np.abs(W1) + abs(cfg.bias)
    kink = (np.abs(z) <= 1e-6 * zc).any(axis=0)
    keep = np.ones(cfg.P, bool)
    keep[:D * H] = np.tile(~kink, D)
    # (the chance that SOME episode puts a unit within 1e-6 of its kink grows with the number of rows)
    assert kink.sum() <= max(2, H // 50, x.shape[0] // 1500), f"{what}: {kink.sum()} hidden units on their ReLU kink"
    g_cmp, ref_cmp = np.where(keep, g, 0.0), np.where(keep, ref["grad"], 0.0)
    if n_terms >= 2048:
        # Thousands of episodes: the oracle adds the episodes' gradients one after the other in fp32 (the device adds
        # in a tree), and with the reference's unscaled N(0,1) weights dq = 2 (q - target) inherits the fp32 rounding of
        # Q itself (|Q| ~ 1e4 from terms of 1e6).  What "fp32-grade" means for such a sum is therefore MEASURED: the
        # fp32 oracle against its fp64 twin (same imposed choices); the device must be within the stated tolerance of
        # the twin, or within twice the fp32 oracle's own distance from it.
        o64 = O.Oracle(cfg, dtype=np.float64, params=w, acc=acc)
        ev = o.env
        o64.env_create(ev["prices"], ev["soe"], ev["B0"], ev["N0"], env_id_base=ev["base"])
        o64.env["budget"][:], o64.env["shares"][:], o64.env["share_value"][:], o64.env["step"] = b, s, v, i
        r64 = o64.env_step(1, trace=True, want_grad=True, force_action=tr["action_sel"], force_knext=tr["k_next"])[0]
        g64 = np.where(keep, r64["grad"], 0.0)
        nrm, m64 = np.linalg.norm(g64), np.abs(g64).max()
        e32_fro = np.linalg.norm(ref_cmp - g64) / nrm
        e32_max = np.abs(ref_cmp - g64).max()
        dev_fro = np.linalg.norm(g_cmp - g64) / nrm
        dev_max = np.abs(g_cmp - g64).max()
        assert dev_fro <= max(_tol.GRAD_FRO, 2 * e32_fro), \
            f"{what} grad vs fp64 twin: relative Frobenius error {dev_fro:.3g} (fp32 oracle: {e32_fro:.3g})"
        assert dev_max <= max(_tol.GRAD_RTOL * m64, 2 * e32_max), \
            f"{what} grad vs fp64 twin: max |dg| {dev_max:.3g} (fp32 oracle: {e32_max:.3g}, max |g| {m64:.3g})"
        if tally is not None:
            tally.grad_err_vs_twin.append((float(dev_fro), float(e32_fro)))
    else:
        _tol.assert_grad_close(g_cmp, ref_cmp, what + " grad", n_terms=n_terms)
    p, a = pol.get_weights()
    # (1) the device's AdaGrad step applied to the device's own gradient sum: the oracle's arithmetic bit for bit
    chk = O.Oracle(cfg, params=w, acc=acc)
    chk.adagrad(g)
    np.testing.assert_array_equal(p, chk.params, err_msg=what + " AdaGrad(params) is not the oracle's arithmetic")
    np.testing.assert_array_equal(a, chk.acc, err_msg=what + " AdaGrad(acc) is not the oracle's arithmetic")
    # (2) against the oracle's weights, where the sign of the gradient element is beyond doubt (with the
    # reference's 1e-8 initial accumulator the first step is lr * sign(g), whatever the magnitude of g)
    m = np.abs(ref["grad"]).max()
    stable = (np.abs(ref["grad"]) > 10 * _tol.GRAD_RTOL * m) & keep
    if stable.any():
        assert np.abs(p[stable] - o.params[stable]).max() <= _tol.W_ATOL, what + " weights"
    if check_state:
        bd, sd, vd = pol.env_state()
        np.testing.assert_array_equal(bd, o.env["budget"], err_msg=what + " carried budget")
        np.testing.assert_array_equal(sd, o.env["shares"], err_msg=what + " carried shares")
        np.testing.assert_array_equal(vd, o.env["share_value"], err_msg=what + " carried share value")
    if tally is not None:
        tally.steps += 1
        tally.rows += len(soe)
        tally.forced_actions += int(n_act)
        tally.forced_knext += int(n_k)
        tally.grad_checked += 1
        tally.kink_columns += int(kink.sum())
    return tr, ref
