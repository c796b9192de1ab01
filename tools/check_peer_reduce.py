"""Run under torchrun on >= 2 GPUs: the peer-memory reduce against the NCCL reduce and the oracle,
then timings of both.  Development / verification tool (needs more than one GPU, so it is not in
the pytest suite that runs on a single B200)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import feedforward as ff                                 # noqa: E402
from rafiki_b200 import engine, synth                                # noqa: E402
from rafiki_b200.distributed import DistributedEnsemble, PeerGather, PeerReducer, trials_of_rank   # noqa: E402


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    W, n = 8, 65536
    dead = synth.dead_mask(784)
    params = [synth.make_trial(t, 28, 1, 2, 128, dead=dead) for t in range(W)]
    mine = trials_of_rank(W, world, rank)
    trials = [engine.Trial(params[t]) for t in mine]
    ens = engine.Ensemble(trials, total_trials=W) if trials else None
    dens = DistributedEnsemble(ens, W, 10, dev)
    red = PeerGather(n, 10, dev) if os.environ.get('RFK_PEER_MODE', 'gather') == 'gather' else PeerReducer(n, 10, dev)
    xs = [torch.from_numpy(synth.make_queries(n, 28, 1, 'uniform', seed=i, dead=dead)).to(dev) for i in range(3)]
    out_nccl = torch.zeros((n, 10), dtype=torch.float32, device=dev)
    ok = True
    for i in range(5):
        x = xs[i % 3]
        res = dens.forward_peer(x, red)
        red.wait()
        dens.forward(x, out=out_nccl)
        torch.cuda.synchronize()
        if rank == 0:
            a, b = res.cpu().numpy(), out_nccl.cpu().numpy()
            sl = slice(1000, 1512)
            want = ff.ensemble_mean_array([ff.forward(p, xs[i % 3][sl].cpu().numpy(), np.float32) for p in params])
            e1, e2 = np.abs(a - b).max(), np.abs(a[sl] - want).max()
            print('step', i, 'peer vs nccl max diff %.3e, peer vs oracle %.3e' % (e1, e2), flush=True)
            ok = ok and e1 <= 1e-6 and e2 <= 2e-3
    for name, fn in (('peer', lambda x: dens.forward_peer(x, red)), ('nccl', lambda x: dens.forward(x, out=out_nccl))):
        for i in range(10):
            fn(xs[i % 3])
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(200):
            fn(xs[i % 3])
        red.wait(0)
        red.wait(1)
        e1.record()
        dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 200], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print('%s reduce: %.1f us per step, %.0f Mq/s' % (name, t.item() * 1e3, n / t.item() / 1e3), flush=True)
    if rank == 0:
        print('PEER CHECK', 'OK' if ok else 'FAILED', flush=True)
    dist.barrier()
    dist.destroy_process_group()


main()
