"""The N > 1 host logic on CPU: trial placement and the reduce of partial means over a
world_size-2 and world_size-3 gloo group.  The per-rank engine is a CPU stand-in built on the
oracle (tests may use it); the product's per-rank engine is the CUDA ``Ensemble``."""
import os
import socket
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import feedforward as ff
from rafiki_b200 import synth
from rafiki_b200.distributed import DistributedEnsemble, trials_of_rank


def test_trial_placement():
    assert trials_of_rank(8, 8, 3) == [3]
    assert trials_of_rank(8, 2, 1) == [1, 3, 5, 7]
    assert trials_of_rank(3, 4, 3) == []
    assert sorted(sum((trials_of_rank(5, 3, r) for r in range(3)), [])) == list(range(5))
    with pytest.raises(ValueError):
        trials_of_rank(4, 2, 2)


class _OraclePartial:
    """What a rank's Ensemble(total_trials=W) returns: sum over its trials of p_t / W."""

    def __init__(self, params, total):
        self.params, self.total = params, total

    def forward(self, x, out):
        acc = np.zeros((x.shape[0], 10))
        for p in self.params:
            acc += ff.forward(p, x.numpy(), np.float32).astype(np.float64)
        out.copy_(torch.from_numpy((acc / self.total).astype(np.float32)))
        return out


def _worker(rank, world, port, n_trials, path):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        dead = synth.dead_mask(784)
        params = [synth.make_trial(t, 28, 1, 1 + t % 2, 8 + 5 * t, dead=dead) for t in range(n_trials)]
        X = synth.make_queries(50, 28, 1, 'uniform', seed=3, dead=dead)
        mine = trials_of_rank(n_trials, world, rank)
        local = _OraclePartial([params[t] for t in mine], n_trials) if mine else None
        ens = DistributedEnsemble(local, n_trials, 10, 'cpu', root=0)
        out = ens.forward(torch.from_numpy(X))
        if rank == 0:
            want = ff.ensemble_mean_array([ff.forward(p, X, np.float32) for p in params])
            np.save(path, np.abs(out.numpy().astype(np.float64) - want).max())
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


@pytest.mark.parametrize('world,n_trials', [(2, 3), (3, 2), (2, 8)])
def test_reduce_of_partial_means_matches_the_full_ensemble(world, n_trials):
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, 'err.npy')
        mp.spawn(_worker, args=(world, _free_port(), n_trials, path), nprocs=world, join=True)
        assert float(np.load(path)) <= 1e-6
