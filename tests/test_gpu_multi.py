"""Multi-GPU parity on the GPU box: the trial-per-GPU ensemble (one process per GPU, peer-memory
all-gather of the queries and reduce of the probabilities over NVLink) must equal the single-GPU
ensemble of the same batch.  Needs >= 2 visible GPUs; on a single-GPU box the test is skipped (the
N > 1 host logic is then covered by the gloo tests in test_distributed_cpu.py, and bench.py checks
`parity_vs_single_gpu` in every N > 1 run)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs at least 2 GPUs')
@pytest.mark.parametrize('n_trials', [5, 2, 3])   # 3 + 2 trials per rank (scatter-mean kernel); one per rank (hand-shake inside the trial kernel); 2 + 1 (both)
def test_peer_ensemble_equals_the_single_gpu_ensemble(n_trials):
    world = 2
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world),
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(ROOT, 'tools', 'check_peer_ensemble.py'), '20000', str(n_trials)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and 'PEER ENSEMBLE OK' in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
