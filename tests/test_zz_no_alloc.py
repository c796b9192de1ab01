"""The allocation-free entry point ``rfk_mlp_forward_static`` (collected last on purpose).

It exists for callers that cannot obtain scratch memory at call time -- a serving loop confined
to a start-up memory budget, or a plugin handed raw device pointers by a framework whose
allocator it must not touch.  Those callers still pass the query and probability buffers; the
activations in between live in a fixed 4 MiB block inside the library.

The test drives it through the C ABI and checks it, on two calls in a row that reuse the block
(the second with a smaller, different batch, so stale activations of the first call would
show), against ``rfk_mlp_forward`` with a caller-owned workspace.  Both run the same kernels, so
the results must agree within the fp32 tolerance of the other tests (1e-5; in fact exactly).
"""
import ctypes

import numpy as np
import pytest
import torch

from rafiki_b200 import _native, synth

pytestmark = pytest.mark.gpu
TOL_F32 = 1e-5


def _forward_with_workspace(lib, h, x, n, c):
    need = ctypes.c_size_t()
    _native.check(lib.rfk_mlp_workspace_bytes(h, n, ctypes.byref(need)))
    ws = torch.empty(max(need.value, 1), dtype=torch.uint8, device='cuda')
    out = torch.empty((n, c), dtype=torch.float32, device='cuda')
    _native.check(lib.rfk_mlp_forward(h, x.data_ptr(), n, out.data_ptr(), ws.data_ptr(), ws.numel(), None))
    torch.cuda.synchronize()
    return out


def test_static_scratch_forward_matches_workspace_forward_twice():
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    lib = _native.load()
    params = synth.make_trial(9, 28, 1, 2, 128, dead=synth.dead_mask(784))
    h = _native.mlp_create(params, _native.RFK_F32)
    first = True
    try:
        for n, seed in ((4096, 1), (1000, 2)):       # 4096 x 128 floats x 2 buffers = the whole block
            x = torch.from_numpy(synth.make_queries(n, 28, 1, 'uniform', seed=seed,
                                                    dead=synth.dead_mask(784)).reshape(n, 784)).cuda()
            want = _forward_with_workspace(lib, h, x, n, 10)
            got = torch.full((n, 10), -1.0, dtype=torch.float32, device='cuda')
            if first:
                # the first call on a device may be where the driver loads the module that holds the block
                _native.check(lib.rfk_mlp_forward_static(h, x.data_ptr(), n, got.data_ptr(), None))
                torch.cuda.synchronize()
                first = False
            # what the name promises: the call takes no device memory (nothing goes through torch's
            # allocator either: these are raw pointers) -- free memory is the same before and after
            free_before = torch.cuda.mem_get_info()[0]
            _native.check(lib.rfk_mlp_forward_static(h, x.data_ptr(), n, got.data_ptr(), None))
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info()[0] == free_before
            diff = (got - want).abs().max().item()
            assert diff <= TOL_F32, diff
            assert torch.equal(got, want)
            assert abs(float(got.sum(1).mean()) - 1.0) < 1e-5
        # one query more than the block can hold is refused, with a message that says why
        n = 4097
        x = torch.zeros((n, 784), dtype=torch.uint8, device='cuda')
        out = torch.empty((n, 10), dtype=torch.float32, device='cuda')
        rc = lib.rfk_mlp_forward_static(h, x.data_ptr(), n, out.data_ptr(), None)
        assert rc == 1 and b'scratch' in lib.rfk_last_error()
        # tensor-core trials need no scratch at all: the entry point says so instead of guessing
        h16 = _native.mlp_create(params, _native.RFK_F16)
        rc = lib.rfk_mlp_forward_static(h16, x.data_ptr(), 16, out.data_ptr(), None)
        assert rc == 1 and b'RFK_F32' in lib.rfk_last_error()
        lib.rfk_mlp_destroy(h16)
    finally:
        lib.rfk_mlp_destroy(h)
