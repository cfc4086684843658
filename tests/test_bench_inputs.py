"""bench.py's synthetic workloads (CPU): the shards the ranks generate are the slices of the one-GPU workload, for the
default workload (one series per episode) and for the router fan-out (256 episodes per series)."""
import importlib
import types

import numpy as np

from oracle import oracle as O


def _bench(per_series):
    b = importlib.import_module("bench")
    b.EPISODES_PER_SERIES = per_series
    return b


def test_default_workload_shards_are_slices():
    b = _bench(1)
    try:
        prices, soe = b.workload_inputs(O, 0, 96)
        assert prices.shape == (96, b.SERIES_LEN) and (soe == np.arange(96)).all()
        for rank in range(4):
            p, s = b.workload_inputs(O, rank * 24, 24)
            np.testing.assert_array_equal(p, prices[rank * 24:(rank + 1) * 24])
            assert (s == np.arange(24)).all()
    finally:
        b.EPISODES_PER_SERIES = 1


def test_fanout_workload_groups_episodes_by_series():
    b = _bench(256)
    try:
        prices, soe = b.workload_inputs(O, 0, 1024)
        assert prices.shape == (4, b.SERIES_LEN)
        assert (soe == np.repeat(np.arange(4), 256)).all() and (np.diff(soe) >= 0).all()
        for rank in range(2):                                   # two ranks: 512 episodes = 2 series each
            p, s = b.workload_inputs(O, rank * 512, 512)
            np.testing.assert_array_equal(p, prices[rank * 2:(rank + 1) * 2])
            assert (s == np.repeat(np.arange(2), 256)).all()
        p, s = b.workload_inputs(O, 384, 256)                   # a shard that starts inside a series
        assert p.shape[0] == 2 and (s == np.repeat(np.arange(2), 128)).all()
        np.testing.assert_array_equal(p, prices[1:3])
    finally:
        b.EPISODES_PER_SERIES = 1


def test_set_workload_fanout_defaults():
    b = importlib.import_module("bench")
    args = types.SimpleNamespace(envs_per_gpu=4096, rollout_steps=0, gpus=1)
    saved = (b.WORKLOAD_NAME, b.EPISODES_PER_SERIES, b.ROLLOUT)
    try:
        b.set_workload("fanout", args)
        assert args.envs_per_gpu == 131072 and b.EPISODES_PER_SERIES == 256 and "configs[4]" in b.WORKLOAD_NAME
    finally:
        b.WORKLOAD_NAME, b.EPISODES_PER_SERIES, b.ROLLOUT = saved
