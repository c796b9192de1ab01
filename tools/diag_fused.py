"""GPU diagnostic for the fused tcgen05 kernel: structured weights that expose which input
element / hidden unit each output saw, then random trials of growing size, then a timing.
Writes gpurun_out/diag_fused.txt.  Development tool, not part of the test suite."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import feedforward as ff           # noqa: E402
from rafiki_b200 import engine, synth          # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'gpurun_out')
os.makedirs(OUT, exist_ok=True)
log = open(os.path.join(OUT, 'diag_fused.txt'), 'w')


def say(*a):
    s = ' '.join(str(x) for x in a)
    print(s, flush=True)
    log.write(s + '\n')
    log.flush()


def identity_bn(d):
    return dict(bn_gamma=np.ones(d, np.float32), bn_beta=np.zeros(d, np.float32), bn_mean=np.zeros(d, np.float32),
                bn_var=np.full(d, 1.0 - 1e-3, np.float32), bn_eps=1e-3)


def probe_trial(d, H, ks, L=1, s=1.0 / 64):
    """hidden_j = x[ks[j]] for j < len(ks) (<= 10); logit_c = s * hidden_c."""
    p = identity_bn(d)
    k1 = np.zeros((d, H), np.float32)
    for j, k in enumerate(ks):
        k1[k, j] = 1.0
    kernels, biases = [k1], [np.zeros(H, np.float32)]
    if L == 2:
        kernels.append(np.eye(H, dtype=np.float32))
        biases.append(np.zeros(H, np.float32))
    ko = np.zeros((H, 10), np.float32)
    for c in range(min(10, H)):
        ko[c, c] = s
    kernels.append(ko)
    biases.append(np.zeros(10, np.float32))
    p.update(image_size=int(round(d ** 0.5)), kernels=kernels, biases=biases)
    return p


def run(params, X, compute='f16'):
    tr = engine.Trial(params, compute=compute)
    out = tr.forward(torch.from_numpy(X).cuda())
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    tr.close()
    return o


def main():
    torch.cuda.set_device(0)
    say('device', torch.cuda.get_device_name(0))
    rng = np.random.default_rng(0)
    d = 784
    # 1. structured probes
    for (H, L, n) in [(16, 1, 128), (16, 2, 128), (128, 1, 128), (128, 2, 300), (48, 2, 1000)]:
        for ks in ([1, 2, 3, 4, 5, 6, 7, 8, 9, 10], [0, 15, 16, 63, 64, 65, 127, 128, 700, 783],
                   [768, 769, 770, 775, 776, 780, 781, 782, 783, 500]):
            X = rng.integers(0, 256, size=(n, d), dtype=np.uint8)
            p = probe_trial(d, H, ks, L)
            try:
                got = run(p, X)
            except Exception as e:          # noqa: BLE001
                say('PROBE H', H, 'L', L, 'n', n, 'ks', ks, 'EXC', repr(e))
                return 1
            want = ff.forward(p, X, np.float64)
            err = np.abs(got - want)
            say('PROBE H', H, 'L', L, 'n', n, 'ks', ks[:4], 'maxerr %.3e' % err.max(), 'bad rows', int((err.max(1) > 2e-3).sum()))
            if err.max() > 2e-3:
                r = int(err.max(1).argmax())
                lg = np.log(np.maximum(got[r], 1e-30))
                lw = np.log(want[r])
                say('  row', r, 'x[ks]', X[r, ks].tolist())
                say('  got dlogit*64 ', np.round((lg - lg[0]) * 64, 1).tolist())
                say('  want dlogit*64', np.round((lw - lw[0]) * 64, 1).tolist())
                bad_rows = np.nonzero(err.max(1) > 2e-3)[0]
                say('  bad rows (first 40):', bad_rows[:40].tolist())
    # 2. random trials
    for (S, ch, L, H, n) in [(28, 1, 1, 16, 128), (28, 1, 2, 128, 128), (28, 1, 2, 128, 4096), (28, 1, 1, 36, 517),
                             (32, 1, 2, 100, 1000), (32, 3, 2, 64, 300), (28, 1, 2, 128, 65536)]:
        dd = S * S * ch
        dead = synth.dead_mask(dd)
        p = synth.make_trial(1, S, ch, L, H, dead=dead, data_kind='fashion')
        X = synth.make_queries(n, S, ch, 'fashion', seed=1, dead=dead)
        try:
            got = run(p, X)
        except Exception as e:          # noqa: BLE001
            say('RAND', S, ch, L, H, n, 'EXC', repr(e))
            return 1
        want = ff.forward(p, X[:8192], np.float32)
        err = np.abs(got[:8192] - want)
        gb = run(p, X, 'bf16')
        say('RAND S', S, 'ch', ch, 'L', L, 'H', H, 'n', n, 'f16 maxerr %.3e' % err.max(),
            'argmax agree %.5f' % np.mean(got[:8192].argmax(1) == want.argmax(1)),
            '| bf16 maxerr %.3e' % np.abs(gb[:8192] - want).max(),
            'argmax agree %.5f' % np.mean(gb[:8192].argmax(1) == want.argmax(1)))
    # 3. timing (device resident)
    dead = synth.dead_mask(784)
    for w, n in [(1, 65536), (3, 4096), (3, 65536), (8, 65536)]:
        trials = [engine.Trial(synth.make_trial(t, 28, 1, 2, 128, dead=dead)) for t in range(w)]
        ens = engine.Ensemble(trials)
        xs = [torch.from_numpy(synth.make_queries(n, 28, 1, 'uniform', seed=i, dead=dead)).cuda() for i in range(4)]
        out = torch.empty((n, 10), dtype=torch.float32, device='cuda')
        for i in range(5):
            ens.forward(xs[i % 4], out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 50
        e0.record()
        for i in range(iters):
            ens.forward(xs[i % 4], out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        flops = n * sum(synth.trial_flops_per_query(synth.make_trial(t, 28, 1, 2, 128)) for t in range(w))
        say('TIME w', w, 'n', n, '%.1f us/launch' % (ms * 1e3), '%.1f Mq/s' % (n / ms / 1e3),
            '%.1f TFLOP/s' % (flops / ms / 1e9), '%.1f GB/s' % (n * 824 / ms / 1e6))
    say('DIAG DONE')
    return 0


if __name__ == '__main__':
    sys.exit(main())
