"""Phase stamps of episode_tc_kernel with the gradient exchange over NVLink peer memory:
torchrun --nproc-per-node 2 profiles/ep_tc_phases_dist.py (rank 0 prints)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import oracle as O

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("gloo", rank=rank, world_size=world)
n, steps = 4096, 60
first, count = q.shard(n * world, rank, world)
prices = O.gen_prices(42, first, count, 1000)
with q.QDecisionPolicy(q.PolicyConfig(device=rank)) as pol:
    pol.set_option("phase_clocks", 1)
    pol.dist_init(dist, rank, world, p2p=True)
    pol.env_create(prices, np.arange(count, dtype=np.int32), 1000.0, 0, env_id_base=first)
    assert pol.get_option("ep_tc_active") and pol.get_option("p2p_active")
    pol.env_step(steps)
    dist.barrier()
    pol.env_step(steps)
    c = pol.phase_clocks(steps)[5:]
if rank == 0:
    f = lambda a, b: (c[:, a] - c[:, b]).mean()
    tot = (c[1:, 0] - c[:-1, 0]).mean()
    print(f"step total {tot:.0f} cyc; phase A {f(8, 0):.0f}; grid.sync {f(9, 8):.0f}; local sums {f(15, 9):.0f}; "
          f"exchange (push, flag, wait, add) {f(12, 15):.0f}; AdaGrad + publish {f(10, 12):.0f}")
dist.barrier()
dist.destroy_process_group()
