"""Achieved HBM bandwidth of the memory-bound kernels (decode / resize / ensemble mean), CUDA events,
inputs larger than L2 rotated between launches.  Prints one JSON line (development / profiles tool)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rafiki_b200 import engine  # noqa: E402


def timed(fn, iters=50):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(iters):
        fn(i)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e-3


def main():
    torch.cuda.set_device(0)
    peak = 6542.1
    try:
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')))['hbm_gbs']
    except Exception:
        pass
    out = {'peak_hbm_gbs': peak, 'kernels': []}
    n = 262144                                             # 4 x the headline batch: 205 MB of uint8 per buffer
    xs = [torch.randint(0, 256, (n, 784), dtype=torch.uint8, device='cuda') for _ in range(2)]
    scale = torch.rand(784, device='cuda') + 0.5
    shift = torch.randn(784, device='cuda')
    for dt, sz in ((torch.bfloat16, 2), (torch.float32, 4)):
        t = timed(lambda i: engine.decode_u8(xs[i & 1], scale, shift, out_dtype=dt))
        byts = n * 784 * (1 + sz)
        out['kernels'].append({'kernel': 'decode_u8 -> %s' % str(dt).split('.')[-1], 'n': n, 'd': 784, 'us': t * 1e6,
                               'algorithmic_bytes': byts, 'gbs': byts / t / 1e9, 'frac': byts / t / 1e9 / peak})
    x28 = [x.view(n, 28, 28) for x in xs]
    t = timed(lambda i: engine.resize_decode_u8(x28[i & 1], 32))
    byts = n * (784 + 1024)
    out['kernels'].append({'kernel': 'resize 28->32 (uint8)', 'n': n, 'us': t * 1e6, 'algorithmic_bytes': byts,
                           'gbs': byts / t / 1e9, 'frac': byts / t / 1e9 / peak})
    del xs, x28
    for w, rows in ((3, 4 * 1048576), (8, 2 * 1048576)):
        sets = [[torch.rand(rows, 10, device='cuda') for _ in range(w)] for _ in range(2)]
        dst = torch.empty(rows, 10, device='cuda')
        t = timed(lambda i: engine.ensemble_mean(sets[i & 1], out=dst))
        byts = rows * 10 * 4 * (w + 1)
        out['kernels'].append({'kernel': 'ensemble_mean W=%d' % w, 'n': rows, 'us': t * 1e6, 'algorithmic_bytes': byts,
                               'gbs': byts / t / 1e9, 'frac': byts / t / 1e9 / peak})
        del sets, dst
    print(json.dumps(out))


main()
