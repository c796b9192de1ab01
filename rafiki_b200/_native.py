"""ctypes binding of ``librafiki_b200.so`` (declared in ``include/rafiki_b200.h``).

The library is built in-tree by ``__graft_entry__.build()`` (``make -C rafiki_b200/csrc``).
There is no CPU fallback: if the shared object is missing, or a call fails, an exception
is raised.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('RFK_LIB', os.path.join(_HERE, 'librafiki_b200.so'))   # RFK_LIB: development override

RFK_F32 = 0
RFK_BF16 = 1
RFK_F16 = 2
RFK_MAX_TRIALS = 32
RFK_MAX_CLASSES = 16
RFK_STATIC_SCRATCH_BYTES = 4 * 1024 * 1024

_ERR_NAMES = {1: 'RFK_ERR_INVALID', 2: 'RFK_ERR_CUDA', 3: 'RFK_ERR_UNSUPPORTED', 4: 'RFK_ERR_COMM'}


class RafikiNativeError(RuntimeError):
    def __init__(self, code, message):
        super().__init__('{} ({}): {}'.format(_ERR_NAMES.get(code, 'RFK_ERR'), code, message))
        self.code = code


_c = ctypes
_vp, _i, _i64, _sz, _f = _c.c_void_p, _c.c_int, _c.c_int64, _c.c_size_t, _c.c_float

# name -> (restype, argtypes); every symbol include/rafiki_b200.h declares
PROTOTYPES = {
    'rfk_version': (_i, []),
    'rfk_last_error': (_c.c_char_p, []),
    'rfk_launch_count': (_c.c_ulonglong, [_c.c_char_p]),
    'rfk_device_check': (_i, [_i]),
    'rfk_decode_u8': (_i, [_vp, _vp, _i64, _i, _i, _vp, _vp, _i, _vp]),
    'rfk_resize_decode_u8': (_i, [_vp, _vp, _i64, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    'rfk_mlp_create': (_i, [_c.POINTER(_vp), _i, _i, _c.POINTER(_i), _i, _c.POINTER(_vp), _c.POINTER(_vp),
                            _vp, _vp, _vp, _vp, _f, _i]),
    'rfk_mlp_destroy': (_i, [_vp]),
    'rfk_mlp_info': (_i, [_vp, _c.POINTER(_i)]),
    'rfk_mlp_workspace_bytes': (_i, [_vp, _i64, _c.POINTER(_sz)]),
    'rfk_mlp_forward': (_i, [_vp, _vp, _i64, _vp, _vp, _sz, _vp]),
    'rfk_mlp_forward_static': (_i, [_vp, _vp, _i64, _vp, _vp]),
    'rfk_ensemble_mean': (_i, [_c.POINTER(_vp), _i, _i64, _i, _vp, _vp]),
    'rfk_ensemble_sum': (_i, [_c.POINTER(_vp), _i, _i64, _i, _vp, _vp]),
    'rfk_ensemble_workspace_bytes': (_i, [_c.POINTER(_vp), _i, _i64, _c.POINTER(_sz)]),
    'rfk_ensemble_forward': (_i, [_c.POINTER(_vp), _i, _i, _vp, _i64, _vp, _vp, _sz, _vp]),
    'rfk_peer_alloc': (_i, [_sz, _c.POINTER(_vp), _vp]),
    'rfk_peer_open': (_i, [_vp, _c.POINTER(_vp)]),
    'rfk_peer_close': (_i, [_vp]),
    'rfk_peer_free': (_i, [_vp]),
    'rfk_peer_enable': (_i, [_i]),
    'rfk_peer_reduce': (_i, [_c.POINTER(_vp), _c.POINTER(_vp), _i, _i, _i, _c.c_uint, _i64, _vp, _i, _vp]),
    'rfk_peer_wait': (_i, [_vp, _i, _c.c_uint, _vp]),
    'rfk_peer_signal': (_i, [_c.POINTER(_vp), _i, _c.c_uint, _vp]),
    'rfk_peer_wait_flags': (_i, [_vp, _i, _c.c_uint, _vp]),
    'rfk_peer_block_bytes': (_i, [_i, _i64, _i, _i, _c.POINTER(_sz)]),
    'rfk_peer_slice_rows': (_i, [_i64, _i, _i, _c.POINTER(_i64), _c.POINTER(_i64)]),
    'rfk_peer_create': (_i, [_c.POINTER(_vp), _i, _i, _i, _c.POINTER(_vp), _i64, _i, _i]),
    'rfk_peer_destroy': (_i, [_vp]),
    'rfk_peer_upload_slot': (_i, [_vp, _i64, _c.POINTER(_vp), _c.POINTER(_i64), _c.POINTER(_i64)]),
    'rfk_peer_allgather_u8': (_i, [_vp, _i64, _c.POINTER(_vp), _vp]),
    'rfk_peer_ensemble_step': (_i, [_vp, _c.POINTER(_vp), _i, _i, _vp, _i64, _vp, _sz, _c.POINTER(_vp), _vp]),
    'rfk_peer_ensemble_step_pipelined': (_i, [_vp, _c.POINTER(_vp), _i, _i, _vp, _i64, _vp, _sz, _c.POINTER(_vp), _vp]),
    'rfk_peer_flush': (_i, [_vp, _c.POINTER(_vp), _vp]),
    'rfk_peer_set_inputs_resident': (_i, [_vp, _i]),
}


def launch_count(name_part=None):
    """Kernels launched by the library in this process whose name contains ``name_part`` (all if None)."""
    return int(load().rfk_launch_count(name_part.encode() if name_part else None))

_lib = None


def load():
    """Load the shared library (once) and attach prototypes.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            'rafiki_b200: {} not found -- build it with `python -c "import __graft_entry__ as g; g.build()"` '
            '(or `make -C rafiki_b200/csrc`).  There is no CPU fallback.'.format(LIB_PATH))
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)        # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().rfk_last_error()
        raise RafikiNativeError(rc, msg.decode('utf-8', 'replace') if msg else '')


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def mlp_create(params, compute_dtype):
    """``rfk_mlp_create`` from a trial-parameter dict (see ``rafiki_b200/synth.py``) on the
    current CUDA device.  Returns the opaque handle (int)."""
    lib = load()
    kernels = [_f32(k) for k in params['kernels']]
    biases = [_f32(b) for b in params['biases']]
    d = int(kernels[0].shape[0])
    n_hidden = len(kernels) - 1
    units = (ctypes.c_int * max(n_hidden, 1))(*[int(k.shape[1]) for k in kernels[:-1]])
    n_classes = int(kernels[-1].shape[1])
    kp = (ctypes.c_void_p * len(kernels))(*[k.ctypes.data for k in kernels])
    bp = (ctypes.c_void_p * len(biases))(*[b.ctypes.data for b in biases])
    g, b, m, v = (_f32(params[k]) for k in ('bn_gamma', 'bn_beta', 'bn_mean', 'bn_var'))
    for i, (k, bb) in enumerate(zip(kernels, biases)):
        if k.ndim != 2 or bb.shape != (k.shape[1],):
            raise ValueError('layer {}: kernel {} / bias {} shapes do not match'.format(i, k.shape, bb.shape))
        if i > 0 and k.shape[0] != kernels[i - 1].shape[1]:
            raise ValueError('layer {}: input width {} != previous output {}'.format(i, k.shape[0], kernels[i - 1].shape[1]))
    for a in (g, b, m, v):
        if a.shape != (d,):
            raise ValueError('BatchNormalization vectors must have shape ({},)'.format(d))
    h = ctypes.c_void_p()
    check(lib.rfk_mlp_create(ctypes.byref(h), d, n_hidden, units, n_classes, kp, bp,
                             g.ctypes.data, b.ctypes.data, m.ctypes.data, v.ctypes.data,
                             float(params.get('bn_eps', 1e-3)), int(compute_dtype)))
    return h.value


def mlp_info(handle):
    out = (ctypes.c_int * 6)()
    check(load().rfk_mlp_info(handle, out))
    return {'d': out[0], 'n_hidden': out[1], 'n_classes': out[2], 'dtype': out[3],
            'fused': bool(out[4]), 'max_units': out[5]}


def handle_array(handles):
    return (ctypes.c_void_p * len(handles))(*handles)
