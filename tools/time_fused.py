"""Time the fused kernel (device resident) for a few (w, n); development tool."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rafiki_b200 import engine, synth

def main():
    torch.cuda.set_device(0)
    S = int(os.environ.get('RFK_S', '28'))
    CH = int(os.environ.get('RFK_CH', '1'))
    dead = synth.dead_mask(S * S * CH) if CH == 1 else None
    cases = [(1, 65536), (8, 65536)] if len(sys.argv) < 2 else [tuple(int(v) for v in a.split('x')) for a in sys.argv[1:]]
    for w, n in cases:
        H = int(os.environ.get('RFK_H', '128'))
        trials = [engine.Trial(synth.make_trial(t, S, CH, 2, H, dead=dead)) for t in range(w)]
        ens = engine.Ensemble(trials)
        xs = [torch.from_numpy(synth.make_queries(n, S, CH, 'uniform', seed=i, dead=dead)).cuda() for i in range(4)]
        out = torch.empty((n, 10), dtype=torch.float32, device='cuda')
        for i in range(3):
            ens.forward(xs[i % 4], out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = int(os.environ.get('RFK_ITERS', '20'))
        e0.record()
        for i in range(iters):
            ens.forward(xs[i % 4], out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        print('dbg=%s d=%d w %d n %d: %.1f us/launch, %.0f Mq/s, %.0f TFLOP/s' % (
            os.environ.get('RFK_PAIR_DEBUG', '0'), S * S * CH, w, n, ms * 1e3, n / ms / 1e3, n * w * synth.trial_flops_per_query(synth.make_trial(0, S, CH, 2, H)) / ms / 1e9), flush=True)

main()
