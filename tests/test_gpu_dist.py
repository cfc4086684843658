"""Multi-GPU tests (2, 4 and 8 ranks) of the sharded path through the C ABI: qp_dist_init + the per-step gradient
all-reduce (in-kernel over NVLink peer memory, or NCCL between kernels).  A case is skipped on a box with fewer GPUs
than it needs; run with `gpurun --gpus 2|4|8` (logs of those runs: profiles/)."""
import os
import socket

import numpy as np
import pytest

from oracle import oracle as O
from tests import _tol

pytestmark = pytest.mark.gpu

N_LOCAL, N_STEPS, T = 96, 5, 260


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _worker(rank, world, port, out_dir, p2p, ep_tc, n_local, nvls=1, rsag=-1):
    import torch
    import torch.distributed as dist
    import scala_akka_tensorflow_sharetrade_helper_b200 as q
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)   # host plumbing only: carries the NCCL id
    cfg = O.Config(bookkeeping=O.RLPY)
    first, count = q.shard(n_local * world, rank, world)
    prices = O.gen_prices(cfg.seed, first, count, T)
    w = (O.gen_normal(cfg.seed, 0, cfg.P) * 0.05).astype(np.float32)
    with q.QDecisionPolicy(q.PolicyConfig(bookkeeping=1, device=rank)) as pol:
        pol.set_weights(w)
        pol.set_option("ep_tc", ep_tc)
        pol.set_option("rsag", rsag)
        pol.dist_init(dist, rank, world, p2p=bool(p2p), nvls=bool(nvls))
        pol.env_create(prices, np.arange(count), 1000.0, 1, env_id_base=first)
        p2p_active = pol.get_option("p2p_active")
        ep_tc_active = pol.get_option("ep_tc_active")
        nvls_active = pol.get_option("nvls_active")
        tr0 = pol.env_step_trace(1)
        g_first = pol.last_grad()                         # same weights on every configuration: strict check
        tr1 = pol.env_step_trace(N_STEPS - 1)
        tr = {k: np.concatenate([tr0[k], tr1[k]], axis=0) for k in ("shares", "budget", "action_sel", "q_s")}
        pol.env_step(3)                                   # a second launch: flags carry over between launches
        tr2 = pol.env_step_trace(2)
        p, a = pol.get_weights()
        g = pol.last_grad()
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), params=p, acc=a, grad=g, grad_first=g_first, shares=tr["shares"],
             budget=tr["budget"], action_sel=tr["action_sel"], q_s=tr["q_s"], shares2=tr2["shares"],
             p2p_active=p2p_active, ep_tc_active=ep_tc_active, nvls_active=nvls_active)
    dist.barrier()
    dist.destroy_process_group()


# (world, p2p, ep_tc, episodes per rank, nvls, rsag): rsag = 1 / 0 forces the share exchange to reduce-scatter + all-gather /
# all-to-all, -1 leaves the choice to the library (reduce-scatter + all-gather from eight ranks on)
CASES = [(2, 1, 1, N_LOCAL, 1, -1), (2, 1, 1, N_LOCAL, 0, -1), (2, 1, 1, N_LOCAL, 1, 1), (2, 1, 1, N_LOCAL, 0, 1), (2, 1, 0, N_LOCAL, 1, -1),
         (2, 0, 0, N_LOCAL, 1, -1), (2, 1, 1, 5000, 1, -1),
         (4, 1, 1, N_LOCAL, 1, -1), (4, 1, 1, N_LOCAL, 1, 1), (4, 0, 0, N_LOCAL, 1, -1),
         (8, 1, 1, N_LOCAL, 1, -1), (8, 1, 1, N_LOCAL, 0, -1), (8, 1, 1, N_LOCAL, 1, 0), (8, 1, 0, N_LOCAL, 1, -1), (8, 1, 1, 5000, 1, -1)]


@pytest.mark.parametrize("world,p2p,ep_tc,n_local,nvls,rsag", CASES)
def test_n_gpus_match_one_gpu(tmp_path, world, p2p, ep_tc, n_local, nvls, rsag):
    """world ranks, one per GPU (skipped on boxes with fewer), against one GPU holding every episode.
    p2p = 1: gradient all-reduce inside the persistent kernel over NVLink peer memory (ep_tc = 1: the
    tensor-core kernel, share-by-share exchange; ep_tc = 0: the FFMA kernel, whole-gradient exchange);
    p2p = 0: one fused step kernel per env-step + NCCL all-reduce between kernels.
    nvls = 1: the exchange blocks are torch symmetric memory; where the system has NVLS the pair kernel sends each share as one
    multicast store (the printed nvls_active says whether it did).
    n_local = 5000: more than one tile per SM, the tensor-core kernel runs in waves on both sides."""
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    import scala_akka_tensorflow_sharetrade_helper_b200 as q
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, world, port, str(tmp_path), p2p, ep_tc, n_local, nvls, rsag)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    parts = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    assert all(int(p["p2p_active"]) == p2p for p in parts), "peer-memory path not taken as requested"
    assert all(int(p["ep_tc_active"]) == ep_tc for p in parts), "tensor-core kernel not (de)selected as requested"
    assert len({int(p["nvls_active"]) for p in parts}) == 1, "ranks disagree on the multicast exchange"
    if not nvls:
        assert int(parts[0]["nvls_active"]) == 0
    print(f"world {world} p2p {p2p} ep_tc {ep_tc} n_local {n_local}: nvls_active = {int(parts[0]['nvls_active'])}")
    # replicas bit-identical: same all-reduced sum, same AdaGrad arithmetic
    for r in range(1, world):
        np.testing.assert_array_equal(parts[0]["params"], parts[r]["params"], err_msg=f"rank {r} params")
        np.testing.assert_array_equal(parts[0]["acc"], parts[r]["acc"], err_msg=f"rank {r} acc")
        np.testing.assert_array_equal(parts[0]["grad"], parts[r]["grad"], err_msg=f"rank {r} grad")
    # against one GPU holding every episode
    cfg = O.Config(bookkeeping=O.RLPY)
    n = n_local * world
    prices = O.gen_prices(cfg.seed, 0, n, T)
    w = (O.gen_normal(cfg.seed, 0, cfg.P) * 0.05).astype(np.float32)
    with q.QDecisionPolicy(q.PolicyConfig(bookkeeping=1)) as pol:
        pol.set_weights(w)
        pol.env_create(prices, np.arange(n), 1000.0, 1)
        tr0 = pol.env_step_trace(1)
        g1_first = pol.last_grad()
        tr1 = pol.env_step_trace(N_STEPS - 1)
        tr = {k: np.concatenate([tr0[k], tr1[k]], axis=0) for k in ("shares", "budget", "action_sel")}
        pol.env_step(3)
        tr2 = pol.env_step_trace(2)
        p1, _ = pol.get_weights()
        g1 = pol.last_grad()
    for k in ("shares", "budget", "action_sel"):
        np.testing.assert_array_equal(np.concatenate([p[k] for p in parts], axis=1), tr[k], err_msg=k)
    np.testing.assert_array_equal(np.concatenate([p["shares2"] for p in parts], axis=1), tr2["shares"])
    _tol.assert_grad_close(parts[0]["grad_first"], g1_first, f"{world}-GPU vs 1-GPU gradient of the first step", n_terms=n)
    # ten updates later the two runs stand at slightly different weights (different summation orders, and the
    # tensor-core kernel on one side only): the last gradients agree to the drift of the weights, not to rounding
    if np.linalg.norm(g1) == 0.0:       # (ten lr = 0.01 steps can leave every ReLU of a small batch dead: then both must be zero)
        assert not parts[0]["grad"].any()
    else:
        d = np.linalg.norm(parts[0]["grad"] - g1) / np.linalg.norm(g1)
        assert d <= 2e-4, f"{world}-GPU vs 1-GPU gradient of the last step: relative Frobenius error {d:.3g}"
    assert np.abs(parts[0]["params"] - p1).max() <= (N_STEPS + 5) * _tol.W_ATOL
