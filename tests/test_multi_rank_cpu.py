"""world_size-2 tests on CPU (gloo) of the multi-GPU algorithm: episodes sharded by rank with global
RNG counters, ONE SUM all-reduce of the gradient per env-step, identical AdaGrad step on every rank.
The arithmetic is the oracle's; what is under test is the sharding + exchange logic that
qp_dist_init / bench.py use on the GPUs (the NCCL path itself is covered by the gpu-marked test)."""
import os
import socket

import numpy as np
import pytest

from oracle import oracle as O
from scala_akka_tensorflow_sharetrade_helper_b200.sharding import shard
from tests import _tol

N_ENVS, N_STEPS, T = 10, 6, 240


def test_shard_covers_everything_once():
    for n, w in ((10, 2), (4096, 8), (7, 3), (1, 1), (5, 8)):
        spans = [shard(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and sum(c for _, c in spans) == n
        for (f0, c0), (f1, _) in zip(spans, spans[1:]):
            assert f0 + c0 == f1
    with pytest.raises(ValueError):
        shard(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = O.Config(bookkeeping=O.RLPY)
    first, count = shard(N_ENVS, rank, world)
    prices = O.gen_prices(cfg.seed, first, count, T)            # this rank's series only
    w = (O.gen_normal(cfg.seed, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, np.arange(count), 1000.0, 1, env_id_base=first)
    rows = []
    for _ in range(N_STEPS):
        p0, a0 = o.params.copy(), o.acc.copy()
        tr = o.env_step(1, trace=True, want_grad=True)[0]        # local gradient (and a local update ...)
        o.params[:], o.acc[:] = p0, a0                           # ... which is undone: the update needs the global sum
        g = torch.from_numpy(tr["grad"].copy())
        dist.all_reduce(g, op=dist.ReduceOp.SUM)                 # the ONE exchange of an env-step
        o.adagrad(g.numpy())
        rows.append(tr)
    # replicas must be bit-identical without any weight broadcast
    mine = torch.from_numpy(o.params.copy())
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    assert all(torch.equal(gathered[0], t) for t in gathered)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), params=o.params, first=first, count=count,
             shares=np.stack([r["shares"] for r in rows]), budget=np.stack([r["budget"] for r in rows]),
             action_sel=np.stack([r["action_sel"] for r in rows]), portfolios=o.portfolios())
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_match_one_process(tmp_path):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    # single-process run over all episodes
    cfg = O.Config(bookkeeping=O.RLPY)
    prices = O.gen_prices(cfg.seed, 0, N_ENVS, T)
    w = (O.gen_normal(cfg.seed, 0, cfg.P) * 0.05).astype(np.float32)
    o = O.Oracle(cfg, params=w)
    o.env_create(prices, np.arange(N_ENVS), 1000.0, 1)
    tr = o.env_step(N_STEPS, trace=True)
    parts = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    assert [int(p["first"]) for p in parts] == [0, 5]
    for k in ("shares", "budget", "action_sel"):
        whole = np.stack([t[k] for t in tr])
        np.testing.assert_array_equal(np.concatenate([p[k] for p in parts], axis=1), whole, err_msg=k)
    np.testing.assert_array_equal(np.concatenate([p["portfolios"] for p in parts]), o.portfolios())
    # the summation order differs (per-rank partial sums), so weights agree within tolerance only
    assert np.abs(parts[0]["params"] - o.params).max() <= N_STEPS * _tol.W_ATOL
    np.testing.assert_array_equal(parts[0]["params"], parts[1]["params"])
