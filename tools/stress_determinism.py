"""Determinism stress of the fused trial/ensemble kernel (development tool, GPU).

The same (queries, trials) launched many times, and W identical copies of one trial against
that trial alone, must give the same bits every time: the float64 mean of W identical fp32
rows is exact, so any difference is the kernel's own.  Every output is compared bit for bit
with the first launch; mismatching rows are located (tile, row in tile) and sized.

    python tools/stress_determinism.py [--reps 30] [--sizes 65536,20000] [--load]

--load keeps a copy stream busy (device-to-device copies) while the kernels run.
Prints one summary line per case and `STRESS mismatches=<total>`; exit code 1 on any mismatch.
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rafiki_b200 import engine, synth          # noqa: E402


def describe(ref, got):
    bad = (ref != got).any(dim=1)
    rows = torch.nonzero(bad).flatten()
    if rows.numel() == 0:
        return None
    diff = (ref[rows].double() - got[rows].double()).abs().max().item()
    r = rows[:8].tolist()
    tiles = sorted(set((x // 128) for x in rows.tolist()))
    return dict(rows=int(rows.numel()), maxdiff=diff, first=[(x // 128, x % 128) for x in r],
                tiles=len(tiles), first_tiles=tiles[:8])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--reps', type=int, default=30)
    ap.add_argument('--sizes', default='65536,20000')
    ap.add_argument('--load', action='store_true')
    ap.add_argument('--hetero', action='store_true', help='also repeat a heterogeneous 3- and 8-trial ensemble')
    args = ap.parse_args()
    torch.cuda.set_device(0)
    dead = synth.dead_mask(784)
    total = 0
    launches = 0
    side = torch.cuda.Stream()
    junk_a = torch.empty(64 << 20, dtype=torch.uint8, device='cuda')
    junk_b = torch.empty_like(junk_a)
    for n in [int(v) for v in args.sizes.split(',')]:
        X = synth.make_queries(n, 28, 1, 'uniform', seed=n & 1023, dead=dead)
        x = torch.from_numpy(X).cuda()
        for (L, H) in [(2, 128), (1, 36)]:
            p = synth.make_trial(3, 28, 1, L, H, dead=dead)
            tr = engine.Trial(p)
            ref = tr.forward(x).clone()
            torch.cuda.synchronize()
            for w in (1, 2, 4, 8):
                ens = engine.Ensemble([tr] * w)
                worst = None
                nbad = 0
                for rep in range(args.reps):
                    if args.load and rep % 2 == 1:
                        with torch.cuda.stream(side):
                            junk_b.copy_(junk_a)
                    got = ens.forward(x)
                    launches += 1
                    d = describe(ref, got)
                    if d is not None:
                        nbad += 1
                        if worst is None or d['rows'] > worst['rows']:
                            worst = d
                torch.cuda.synchronize()
                total += nbad
                print('n %6d L %d H %3d w %d: %d/%d launches differ from the single-trial reference%s' % (
                    n, L, H, w, nbad, args.reps, '' if worst is None else '  worst: ' + repr(worst)), flush=True)
            tr.close()
        if args.hetero:
            shapes = [(2, 128), (1, 36), (2, 7), (2, 128), (1, 128), (2, 64), (1, 2), (2, 100)]
            for w in (3, 8):
                trials = [engine.Trial(synth.make_trial(t, 28, 1, *shapes[t], dead=dead)) for t in range(w)]
                ens = engine.Ensemble(trials)
                ref = ens.forward(x).clone()
                nbad, worst = 0, None
                for rep in range(args.reps):
                    got = ens.forward(x)
                    launches += 1
                    d = describe(ref, got)
                    if d is not None:
                        nbad += 1
                        worst = d if worst is None or d['rows'] > worst['rows'] else worst
                total += nbad
                print('n %6d hetero w %d: %d/%d launches differ from the first%s' % (
                    n, w, nbad, args.reps, '' if worst is None else '  worst: ' + repr(worst)), flush=True)
                for t in trials:
                    t.close()
    print('STRESS launches=%d mismatches=%d RFK_FUSED_DEBUG=%s RFK_LIB=%s' % (
        launches, total, os.environ.get('RFK_FUSED_DEBUG', '0'), os.path.basename(os.environ.get('RFK_LIB', 'default'))),
        flush=True)
    return 1 if total else 0


if __name__ == '__main__':
    sys.exit(main())
