"""Run under torchrun: where a multi-GPU step's time goes (development).  Per rank: the local fused kernel
alone, the whole peer step (kernel + reduce + wait), and the host time of one step call."""
import os
import sys
import time
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rafiki_b200 import engine, synth                                # noqa: E402
from rafiki_b200.distributed import PeerEnsemble, trials_of_rank     # noqa: E402


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    W, n = int(os.environ.get('RFK_TRIALS', '8')), 65536
    dead = synth.dead_mask(784)
    params = [synth.make_trial(t, 28, 1, 2, 128, dead=dead) for t in range(W)]
    trials = [engine.Trial(params[t]) for t in trials_of_rank(W, world, rank)]
    ens = engine.Ensemble(trials, total_trials=W)
    peer = PeerEnsemble(ens, W, n, 784, 10, dev)
    xs = [torch.from_numpy(synth.make_queries(n, 28, 1, 'uniform', seed=i, dead=dead).reshape(n, 784)).to(dev) for i in range(4)]
    out = torch.empty((n, 10), dtype=torch.float32, device=dev)
    iters = 300

    def timed(fn):
        for i in range(10):
            fn(i)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        host = (time.perf_counter() - t0) / iters * 1e6
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3, host], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t[0].item(), t[1].item()

    if os.environ.get('RFK_RESIDENT'):
        peer.set_inputs_resident(True)
    if os.environ.get('RFK_PEER_DEBUG'):
        c = timed(lambda i: peer.step_pipelined(xs[i % 4]))
        if rank == 0:
            print('RFK_PEER_DEBUG=%s world %d: pipelined peer step %.1f us' % (os.environ['RFK_PEER_DEBUG'], world, c[0]), flush=True)
        os._exit(0)
    a = timed(lambda i: ens.forward(xs[i % 4], out=out))
    b = timed(lambda i: peer.step(xs[i % 4]))
    c = timed(lambda i: peer.step_pipelined(xs[i % 4]))
    peer.flush()
    torch.cuda.synchronize()
    peer.set_inputs_resident(True)
    c2 = timed(lambda i: peer.step_pipelined(xs[i % 4]))
    peer.flush()
    if rank == 0:
        print('world %d: pipelined peer step with resident inputs (kernels overlap) %.1f us' % (world, c2[0]), flush=True)
    torch.cuda.synchronize()
    if rank == 0:
        print('world %d: pipelined peer step %.1f us (host %.1f us per call)' % (world, c[0], c[1]), flush=True)
    if rank == 0:
        print('world %d: local fused forward %.1f us (host %.1f us per call); peer step %.1f us (host %.1f us per call)' % (
            world, a[0], a[1], b[0], b[1]), flush=True)
    peer.close()
    dist.barrier()
    dist.destroy_process_group()


main()
