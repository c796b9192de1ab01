"""Run under torchrun on >= 2 GPUs: the one-call-per-step trial-per-GPU ensemble (PeerEnsemble:
rfk_peer_allgather_u8 + rfk_peer_ensemble_step) against the single-GPU ensemble of the same batch
and the oracle.  Used by tests/test_gpu_multi.py (skipped on a single-GPU box) and by hand:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29531 tools/check_peer_ensemble.py [n]

Prints `PEER ENSEMBLE OK` (rank 0) and exits 0 when every step agrees.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import feedforward as ff                                 # noqa: E402  (checker)
from rafiki_b200 import engine, synth                                # noqa: E402
from rafiki_b200.distributed import PeerEnsemble, slice_rows, trials_of_rank   # noqa: E402


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    n_max = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    # default: 5 trials on 2 ranks (uneven placement, 3 + 2), else 8; argv[2] overrides (W = world: one trial
    # per rank, the path whose hand-shake runs inside the trial kernel)
    W = int(sys.argv[2]) if len(sys.argv) > 2 else (5 if world == 2 else 8)
    dead = synth.dead_mask(784)
    shapes = [(2, 128), (1, 36), (2, 77), (2, 128), (1, 128), (2, 64), (1, 2), (2, 100)]
    params = [synth.make_trial(t, 28, 1, *shapes[t], dead=dead, data_kind='fashion') for t in range(W)]
    mine = trials_of_rank(W, world, rank)
    trials = [engine.Trial(params[t]) for t in mine]
    ens = engine.Ensemble(trials, total_trials=W) if trials else None
    peer = PeerEnsemble(ens, W, n_max, 784, 10, dev, root=0)
    full = engine.Ensemble([engine.Trial(p) for p in params]) if rank == 0 else None
    ok = True
    worst = 0.0
    for stepno, n in enumerate([n_max, n_max, 517, n_max - 16, 1000, n_max]):
        n -= n % 2                                  # n * classes must be a multiple of 4: even n
        X = synth.make_queries(n, 28, 1, 'fashion', seed=40 + stepno, dead=dead).reshape(n, 784)
        h = torch.from_numpy(X).pin_memory()
        lo, hi = slice_rows(n, world, rank)
        nbytes = peer.upload(h)
        assert nbytes == (hi - lo) * 784
        xg = peer.gather(n)
        res = peer.step(xg)
        torch.cuda.synchronize()
        same = bool(torch.equal(xg.cpu(), torch.from_numpy(X)))        # every rank holds the whole batch
        flag = torch.tensor([0 if same else 1], device=dev)
        dist.all_reduce(flag)
        if rank == 0:
            want = full.forward(torch.from_numpy(X).to(dev))
            e_gpu = float((res.double() - want.double()).abs().max().item())
            sl = slice(0, min(n, 512))
            oracle = ff.ensemble_mean_array([ff.forward(p, X[sl], np.float32) for p in params])
            e_or = float(np.abs(res[sl].cpu().numpy().astype(np.float64) - oracle).max())
            worst = max(worst, e_gpu)
            print('step %d n %6d: gathered batch %s on all ranks, vs single GPU %.3e, vs oracle %.3e' % (
                stepno, n, 'identical' if int(flag.item()) == 0 else 'DIFFERS', e_gpu, e_or), flush=True)
            ok = ok and int(flag.item()) == 0 and e_gpu <= 1e-6 and e_or <= 2e-3
        dist.barrier()
    # the pipelined form: the result of a batch comes back one call later (or from flush)
    batches = [synth.make_queries(n_max - 2 * i, 28, 1, 'fashion', seed=60 + i, dead=dead).reshape(-1, 784) for i in range(5)]
    got = []
    for X in batches:
        prev = peer.step_pipelined(torch.from_numpy(X).to(dev))
        if prev is not None:
            got.append(prev.clone())
    got.append(peer.flush().clone())
    torch.cuda.synchronize()
    if rank == 0:
        for i, X in enumerate(batches):
            want = full.forward(torch.from_numpy(X).to(dev))
            e = float((got[i].double() - want.double()).abs().max().item())
            worst = max(worst, e)
            print('pipelined batch %d n %6d: vs single GPU %.3e' % (i, X.shape[0], e), flush=True)
            ok = ok and e <= 1e-6 and got[i].shape == want.shape
    dist.barrier()
    # the same with resident batches and the promise that lets consecutive trial kernels overlap
    resident = [torch.from_numpy(X).to(dev) for X in batches]
    torch.cuda.synchronize()
    dist.barrier()
    peer.set_inputs_resident(True)
    got = []
    for rep in range(3):
        for x in resident:
            prev = peer.step_pipelined(x)
            if prev is not None:
                got.append(prev.clone())
    got.append(peer.flush().clone())
    peer.set_inputs_resident(False)
    torch.cuda.synchronize()
    if rank == 0:
        for i, g in enumerate(got):
            want = full.forward(resident[i % len(resident)])
            e = float((g.double() - want.double()).abs().max().item())
            worst = max(worst, e)
            ok = ok and e <= 1e-6 and g.shape == want.shape
        print('pipelined, resident batches, overlapping kernels: %d results, worst vs single GPU %.3e' % (len(got), worst), flush=True)
    dist.barrier()
    verdict = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(verdict, 0)
    if rank == 0:
        print('PEER ENSEMBLE', 'OK' if ok else 'FAILED', 'max |d| vs single GPU %.3e' % worst, flush=True)
    peer.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0 if int(verdict.item()) == 1 else 1


if __name__ == '__main__':
    sys.exit(main())
