"""GPU test of qp_td_grad_noalloc, the entry point for callers that cannot use the CUDA allocator
(an actor sharing its dispatcher's GPU with other streams, or a process whose device memory belongs
to another runtime's pool).  It is called through the C ABI with caller-owned device pointers and
must give the same summed TD gradient as the allocating entry point qp_update_q, on two calls in a
row that share the library's built-in scratch array.  Kept in its own file, collected last."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O
from tests import _tol
from tests.golden.make_golden import grid_weights

pytestmark = pytest.mark.gpu


def _reference_grad(cfg, w, s, r, s2):
    """The existing, allocating path: UpdateQ on a fresh policy, then the gradient it applied."""
    import scala_akka_tensorflow_sharetrade_helper_b200 as q
    with q.QDecisionPolicy(q.PolicyConfig(window=cfg.wn, hidden=cfg.hidden)) as pol:
        pol.set_weights(w)
        pol.update_q(s, r, s2)
        return pol.last_grad()


def _noalloc_grad(cfg, w, s, r, s2):
    import torch
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    lib = _capi.lib()
    dev = torch.device("cuda", 0)
    d_w, d_s, d_r, d_s2 = (torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(dev) for a in (w, s, r, s2))
    d_g = torch.full((cfg.P,), float("nan"), dtype=torch.float32, device=dev)   # every element must be written
    torch.cuda.synchronize()
    c = _capi.QpConfig()
    lib.qp_config_default(C.byref(c))
    c.window, c.hidden = cfg.wn, cfg.hidden
    st = lib.qp_td_grad_noalloc(0, C.byref(c), d_w.data_ptr(), d_s.data_ptr(), d_r.data_ptr(), d_s2.data_ptr(),
                                s.shape[0], d_g.data_ptr(), None)
    assert st == 0, _capi.last_error()
    return d_g.cpu().numpy()


def _batch(cfg, B, seed):
    rng = np.random.default_rng(seed)
    s = rng.uniform(50, 150, (B, cfg.D)).astype(np.float32)
    s2 = s + rng.uniform(-1, 1, (B, cfg.D)).astype(np.float32)
    r = rng.uniform(-5, 5, B).astype(np.float32)
    return s, r, s2


def test_noalloc_gradient_matches_update_q_twice_in_a_row():
    cfg = O.Config()
    w = grid_weights(42, cfg.P) * np.float32(0.05)
    # second call is smaller than the first, so stale words of the first call surround its data
    for B, seed in ((300, 1), (77, 2)):
        s, r, s2 = _batch(cfg, B, seed)
        g_ref = _reference_grad(cfg, w, s, r, s2)
        assert np.abs(g_ref).max() > 0
        g = _noalloc_grad(cfg, w, s, r, s2)
        assert np.isfinite(g).all(), "some gradient element was never written"
        _tol.assert_grad_close(g, g_ref, f"noalloc B={B}")
        # and both agree with the CPU oracle
        g_or, _ = O.Oracle(cfg, params=w).td_grad(s, r, s2)
        _tol.assert_grad_close(g, g_or, f"noalloc vs oracle B={B}")


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    import torch
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    lib = _capi.lib()
    cfg = O.Config()
    B = 5000
    dev = torch.device("cuda", 0)
    z = torch.zeros(B * cfg.D, dtype=torch.float32, device=dev)
    g = torch.zeros(cfg.P, dtype=torch.float32, device=dev)
    c = _capi.QpConfig()
    lib.qp_config_default(C.byref(c))
    st = lib.qp_td_grad_noalloc(0, C.byref(c), g.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), B, g.data_ptr(), None)
    assert st == _capi.QP_EINVAL
    msg = _capi.last_error()
    assert "4194304" in msg and "scratch" in msg
    # the largest documented batch is accepted
    st = lib.qp_td_grad_noalloc(0, C.byref(c), g.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), 1792, g.data_ptr(), None)
    assert st == 0, _capi.last_error()
    st = lib.qp_td_grad_noalloc(0, C.byref(c), g.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), 1793, g.data_ptr(), None)
    assert st == _capi.QP_EINVAL


def test_noalloc_refuses_a_misaligned_weight_pointer():
    """A caller that carves its buffers out of another runtime's pool may hand over a pointer at an odd offset; the
    kernels fetch W1 in 16-byte granules, so such a pointer is refused up front instead of faulting on the GPU."""
    import torch
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    lib = _capi.lib()
    cfg = O.Config()
    dev = torch.device("cuda", 0)
    pool = torch.zeros(cfg.P + 8, dtype=torch.float32, device=dev)
    z = torch.zeros(4 * cfg.D, dtype=torch.float32, device=dev)
    g = torch.zeros(cfg.P, dtype=torch.float32, device=dev)
    c = _capi.QpConfig()
    lib.qp_config_default(C.byref(c))
    st = lib.qp_td_grad_noalloc(0, C.byref(c), pool.data_ptr() + 4, z.data_ptr(), z.data_ptr(), z.data_ptr(), 4, g.data_ptr(), None)
    assert st == _capi.QP_EINVAL and "16-byte" in _capi.last_error()
    st = lib.qp_td_grad_noalloc(0, C.byref(c), pool.data_ptr() + 16, z.data_ptr(), z.data_ptr(), z.data_ptr(), 4, g.data_ptr(), None)
    assert st == 0, _capi.last_error()
    torch.cuda.synchronize()
