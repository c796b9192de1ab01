"""Small run of the non-tcgen05 kernels for compute-sanitizer (development): decode, resize, ensemble mean,
fp32 MLP (workspace and static entry points), single-rank peer reduce, plus one small fused launch."""
import ctypes
import os
import sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rafiki_b200 import _native, engine, synth     # noqa: E402

torch.cuda.set_device(0)
lib = _native.load()
dead = synth.dead_mask(784)
X = synth.make_queries(777, 28, 1, 'fashion', seed=1, dead=dead)
x = torch.from_numpy(X).cuda()
p = synth.make_trial(1, 28, 1, 2, 64, dead=dead)
scale = torch.rand(784, device='cuda') + 0.5
shift = torch.randn(784, device='cuda')
for dt in (torch.float32, torch.bfloat16):
    engine.decode_u8(x.view(777, 784), scale, shift, out_dtype=dt)
engine.resize_decode_u8(x, 32)
probs = [torch.rand(777, 10, device='cuda') for _ in range(3)]
engine.ensemble_mean(probs)
t32 = engine.Trial(p, compute='f32')
a = t32.forward(x)
out = torch.empty_like(a)
_native.check(lib.rfk_mlp_forward_static(t32.handle, x.data_ptr(), 777, out.data_ptr(), None))
flags = torch.zeros(3, dtype=torch.int32, device='cuda')
part = torch.rand((1000, 10), dtype=torch.float32, device='cuda')
res = torch.zeros_like(part)
_native.check(lib.rfk_peer_enable(0))
_native.check(lib.rfk_peer_reduce((ctypes.c_void_p * 1)(part.data_ptr()), (ctypes.c_void_p * 1)(flags.data_ptr()), 1, 0, 0, 1,
                                  10000, res.data_ptr(), 8, None))
_native.check(lib.rfk_peer_wait(flags.data_ptr(), 1, 1, None))
t16 = engine.Trial(p)
b = t16.forward(x)
torch.cuda.synchronize()
assert torch.equal(out, a) and torch.equal(res, part) and float((a - b).abs().max()) < 2e-3
print('SANITIZE RUN OK')
