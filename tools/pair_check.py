"""Development check of the CTA-pair kernel against the single-CTA kernel (RFK_NO_PAIR=1) and the oracle."""
import os
import sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import feedforward as ff          # noqa: E402
from rafiki_b200 import engine, synth         # noqa: E402

torch.cuda.set_device(0)
dead = synth.dead_mask(784)
shapes = [(2, 128), (1, 36), (2, 77), (2, 128), (1, 128), (2, 64), (1, 2), (2, 100)]
ok = True
for n in (20000, 65536, 9473):
    X = synth.make_queries(n, 28, 1, 'fashion', seed=n % 97, dead=dead)
    x = torch.from_numpy(X).cuda()
    for w in (1, 3, 8):
        params = [synth.make_trial(t, 28, 1, *shapes[t], dead=dead, data_kind='fashion') for t in range(w)]
        trials = [engine.Trial(p) for p in params]
        ens = engine.Ensemble(trials)
        os.environ.pop('RFK_NO_PAIR', None)
        a = ens.forward(x).clone()
        torch.cuda.synchronize()
        os.environ['RFK_NO_PAIR'] = '1'
        b = ens.forward(x).clone()
        torch.cuda.synchronize()
        os.environ.pop('RFK_NO_PAIR', None)
        sl = slice(max(0, n - 700), n)
        want = ff.ensemble_mean_array([ff.forward(p, X[sl], np.float32) for p in params])
        e_or = np.abs(a[sl].cpu().numpy().astype(np.float64) - want).max()
        diff = (a != b).any(dim=1).sum().item()
        md = (a.double() - b.double()).abs().max().item()
        print('n %6d w %d: pair vs single-CTA kernel: %d rows differ (max %.3e); pair vs oracle %.3e' % (n, w, diff, md, e_or), flush=True)
        ok = ok and e_or <= 2e-3 and md <= 1e-6
        for t in trials:
            t.close()
print('PAIR CHECK', 'OK' if ok else 'FAILED')
sys.exit(0 if ok else 1)
