"""Host runtime of the serving path: packed trials, single-GPU ensembles, pinned-host pipeline.

PyTorch is plumbing here (device memory, streams, pinned buffers); every arithmetic step is a
call into ``librafiki_b200.so`` through ``_native``.  Nothing in this module computes on the
CPU and nothing falls back to torch ops: without the CUDA library, construction fails.

Mapping to the reference:

* :class:`Trial`      -- what ``InferenceWorker._load_model`` builds (rafiki/worker/inference.py:78-93):
  one trial's network, ready to answer batches.
* :class:`Ensemble`   -- the ``Predictor`` fan-out/gather/mean (rafiki/predictor/predictor.py:35-70)
  for trials that share a GPU: one fused kernel launch per batch.
* :class:`HostPipeline` -- the Redis hand-off (rafiki/cache/cache.py:36-63) replaced by pinned host
  buffers with async H2D / D2H copies on a copy stream pair.
"""
import ctypes

import numpy as np
import torch

from . import _native

COMPUTE = {'f32': _native.RFK_F32, 'fp32': _native.RFK_F32, 'bf16': _native.RFK_BF16,
           'f16': _native.RFK_F16, 'fp16': _native.RFK_F16}
DEFAULT_COMPUTE = 'f16'


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError('rafiki_b200 needs a CUDA device (sm_100a); there is no CPU fallback')


def _stream_ptr(stream=None):
    s = torch.cuda.current_stream() if stream is None else stream
    return ctypes.c_void_p(s.cuda_stream)


def _check_queries(x, d):
    if not (isinstance(x, torch.Tensor) and x.is_cuda and x.dtype == torch.uint8):
        raise TypeError('queries must be a CUDA uint8 tensor')
    if not x.is_contiguous():
        raise ValueError('queries must be contiguous')
    n = x.shape[0]
    if x.numel() != n * d:
        raise ValueError('queries have {} features per row, the trial expects {}'.format(
            x.numel() // max(n, 1), d))
    return n


class Workspace:
    """Grow-only device scratch for the calls of ONE stream.

    The kernels of a call write their intermediates here (per-trial probabilities of the small-batch
    schedule, activations of the layered paths) and a later kernel of the same call reads them, so
    two calls that may overlap on the GPU (different streams) must not share one: every
    :class:`HostPipeline` slot owns its own, and ``forward(..., workspace=...)`` takes one
    explicitly.  Growing never frees the old block -- kernels already enqueued may still use it --
    it is kept until the object dies.
    """

    def __init__(self, device, nbytes=0):
        self.device = device
        self.buf = None
        self._retired = []
        if nbytes:
            self.get(nbytes)

    def get(self, nbytes):
        if nbytes == 0:
            return ctypes.c_void_p(0), 0
        if self.buf is None or self.buf.numel() < nbytes:
            if self.buf is not None:
                self._retired.append(self.buf)
            self.buf = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return ctypes.c_void_p(self.buf.data_ptr()), self.buf.numel()


_Workspace = Workspace


def fold_resize_into_params(params, s_in, channels=1):
    """Parameters of a trial trained on ``S x S`` images -> equivalent parameters that take the
    ``s_in x s_in`` frames the queries arrive in (rafiki/model/dataset.py:76-78 resizes 28 x 28 queries to
    the ``image_size`` knob, 32 by default, before the network sees them).

    Pillow's NEAREST resize is a gather, and BatchNormalization + the first Dense layer are affine in the
    resized pixels, so both fold into the first layer: ``K'[src] = sum over dst with source src of
    inv[dst] * K[dst]`` and ``b' = b + shift @ K`` (float64 on the host, once).  The kernels then read the
    raw frames: no resize pass over the batch, and a shorter contraction (784 instead of 1024).
    """
    from .model.dataset import nearest_index_table
    S = int(params['image_size'])
    ch = int(channels)
    rows = nearest_index_table(s_in, S)
    cols = nearest_index_table(s_in, S)
    src = ((rows[:, None] * s_in + cols[None, :])[:, :, None] * ch + np.arange(ch)[None, None, :]).reshape(-1)
    eps = float(params.get('bn_eps', 1e-3))
    g, b, m, v = (np.asarray(params[k], np.float64) for k in ('bn_gamma', 'bn_beta', 'bn_mean', 'bn_var'))
    inv = g / np.sqrt(v + eps)
    shift = b - m * inv
    K = np.asarray(params['kernels'][0], np.float64)
    if K.shape[0] != S * S * ch:
        raise ValueError('first kernel has {} rows, expected {}'.format(K.shape[0], S * S * ch))
    d_in = s_in * s_in * ch
    K_eff = np.zeros((d_in, K.shape[1]), np.float64)
    np.add.at(K_eff, src, K * inv[:, None])
    out = dict(params)
    out['kernels'] = [K_eff.astype(np.float32)] + [np.asarray(k, np.float32) for k in params['kernels'][1:]]
    out['biases'] = [(np.asarray(params['biases'][0], np.float64) + shift @ K).astype(np.float32)] + \
        [np.asarray(x, np.float32) for x in params['biases'][1:]]
    out['bn_gamma'] = np.ones(d_in, np.float32)                 # identity BatchNormalization: inv = 1, shift = 0
    out['bn_beta'] = np.zeros(d_in, np.float32)
    out['bn_mean'] = np.zeros(d_in, np.float32)
    out['bn_var'] = np.full(d_in, 1.0 - eps, np.float32)
    out['image_size'] = int(s_in)
    return out


class Trial:
    """One packed trial (TfFeedForward network) on one GPU."""

    def __init__(self, params, compute=DEFAULT_COMPUTE, device=None):
        _require_cuda()
        self._lib = _native.load()
        self.device = torch.device('cuda', torch.cuda.current_device() if device is None else
                                   torch.device(device).index or 0)
        self.compute = compute
        with torch.cuda.device(self.device):
            self.handle = _native.mlp_create(params, COMPUTE[compute])
        info = _native.mlp_info(self.handle)
        self.d = info['d']
        self.n_classes = info['n_classes']
        self.fused = info['fused']
        # geometry of the images the trial was trained on (S x S x ch, flattened to d features)
        self.image_size = int(params['image_size']) if params.get('image_size') else None
        if self.image_size is None or self.d % (self.image_size ** 2):
            self.image_size = next((s for ch in (1, 3) for s in [int(round((self.d / ch) ** 0.5))]
                                    if s * s * ch == self.d), None)
        self.channels = self.d // (self.image_size ** 2) if self.image_size else 1
        self._ws = Workspace(self.device)
        self._params = params
        self._folded = {}

    def for_input(self, s_in):
        """The same trial packed for ``s_in x s_in`` frames (the resize folded into its first layer, see
        :func:`fold_resize_into_params`); tensor-core modes only -- the fp32 parity path keeps the explicit
        resize kernel so that its arithmetic stays the oracle's, operation for operation."""
        s_in = int(s_in)
        if s_in == self.image_size:
            return self
        if self.compute in ('f32', 'fp32'):
            raise ValueError('the fp32 path does not fold the resize')
        t = self._folded.get(s_in)
        if t is None:
            t = self._folded[s_in] = Trial(fold_resize_into_params(self._params, s_in, self.channels),
                                           compute=self.compute, device=self.device)
        return t

    def workspace_bytes(self, n):
        out = ctypes.c_size_t()
        _native.check(self._lib.rfk_mlp_workspace_bytes(self.handle, n, ctypes.byref(out)))
        return out.value

    def forward(self, x, out=None, stream=None, workspace=None):
        """``uint8 [n, ...]`` CUDA tensor -> ``float32 [n, C]`` class probabilities.  ``workspace``:
        a :class:`Workspace` of the calling stream (default: the trial's own, for one stream)."""
        if self.handle is None:
            raise RuntimeError('trial was destroyed')
        n = _check_queries(x, self.d)
        if out is None:
            out = torch.empty((n, self.n_classes), dtype=torch.float32, device=self.device)
        ws, ws_bytes = (workspace or self._ws).get(self.workspace_bytes(n))
        with torch.cuda.device(self.device):
            _native.check(self._lib.rfk_mlp_forward(self.handle, x.data_ptr(), n, out.data_ptr(), ws, ws_bytes,
                                                    _stream_ptr(stream)))
        return out

    def close(self):
        for t in getattr(self, '_folded', {}).values():
            t.close()
        self._folded = {}
        if getattr(self, 'handle', None) is not None:
            self._lib.rfk_mlp_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Ensemble:
    """W trials on one GPU answering the same batch; the mean is taken on the device."""

    def __init__(self, trials, total_trials=None):
        if len(trials) == 0 or len(trials) > _native.RFK_MAX_TRIALS:
            raise ValueError('an ensemble takes 1..{} trials'.format(_native.RFK_MAX_TRIALS))
        self._lib = _native.load()
        self.trials = list(trials)
        self.device = trials[0].device
        self.d = trials[0].d
        self.n_classes = trials[0].n_classes
        self.total_trials = len(trials) if total_trials is None else int(total_trials)
        self._handles = _native.handle_array([t.handle for t in self.trials])
        self._ws = Workspace(self.device)
        self._ws_bytes = {}
        self.fused = all(t.fused for t in self.trials)
        self.image_size = trials[0].image_size
        self.channels = trials[0].channels
        self._folded = {}

    def for_input(self, s_in):
        """The ensemble of the same trials packed for ``s_in x s_in`` frames (see :meth:`Trial.for_input`)."""
        s_in = int(s_in)
        if s_in == self.image_size:
            return self
        e = self._folded.get(s_in)
        if e is None:
            e = self._folded[s_in] = Ensemble([t.for_input(s_in) for t in self.trials], total_trials=self.total_trials)
        return e

    @property
    def can_fold_resize(self):
        return all(t.compute not in ('f32', 'fp32') for t in self.trials)

    def workspace_bytes(self, n):
        out = ctypes.c_size_t()
        _native.check(self._lib.rfk_ensemble_workspace_bytes(self._handles, len(self.trials), n, ctypes.byref(out)))
        return out.value

    def forward(self, x, out=None, stream=None, workspace=None):
        """``workspace``: a :class:`Workspace` of the calling stream (default: the ensemble's own,
        which serves one stream -- calls on several streams each bring theirs)."""
        n = _check_queries(x, self.d)
        if out is None:
            out = torch.empty((n, self.n_classes), dtype=torch.float32, device=self.device)
        need = self._ws_bytes.get(n)
        if need is None:
            need = self._ws_bytes[n] = self.workspace_bytes(n)
        ws, ws_bytes = (workspace or self._ws).get(need)
        if torch.cuda.current_device() == self.device.index:      # the common case: no device switch
            _native.check(self._lib.rfk_ensemble_forward(self._handles, len(self.trials), self.total_trials,
                                                         x.data_ptr(), n, out.data_ptr(), ws, ws_bytes,
                                                         _stream_ptr(stream)))
        else:
            with torch.cuda.device(self.device):
                _native.check(self._lib.rfk_ensemble_forward(self._handles, len(self.trials), self.total_trials,
                                                             x.data_ptr(), n, out.data_ptr(), ws, ws_bytes,
                                                             _stream_ptr(stream)))
        return out


def ensemble_mean(probs, out=None, stream=None):
    """Kernel 3 alone: list of ``float32 [n, C]`` CUDA tensors -> their mean (float64 sum, one rounding)."""
    lib = _native.load()
    w = len(probs)
    n, c = probs[0].shape
    for p in probs:
        if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and tuple(p.shape) == (n, c)):
            raise ValueError('probs must be contiguous float32 CUDA tensors of one shape')
    if out is None:
        out = torch.empty((n, c), dtype=torch.float32, device=probs[0].device)
    ptrs = (ctypes.c_void_p * w)(*[p.data_ptr() for p in probs])
    with torch.cuda.device(out.device):
        _native.check(lib.rfk_ensemble_mean(ptrs, w, n, c, out.data_ptr(), _stream_ptr(stream)))
    return out


def decode_u8(x, scale=None, shift=None, out_dtype=torch.bfloat16, ld_out=None, stream=None):
    """Kernel 1 alone: ``uint8 [n, d]`` -> ``out_dtype [n, ld_out]`` with the BN affine."""
    lib = _native.load()
    n = x.shape[0]
    d = x.numel() // max(n, 1) if n else int(np.prod(x.shape[1:]))
    ld = d if ld_out is None else int(ld_out)
    out = torch.empty((n, ld), dtype=out_dtype, device=x.device)
    code = {torch.float32: _native.RFK_F32, torch.bfloat16: _native.RFK_BF16}[out_dtype]
    sp = ctypes.c_void_p(scale.data_ptr()) if scale is not None else ctypes.c_void_p(0)
    hp = ctypes.c_void_p(shift.data_ptr()) if shift is not None else ctypes.c_void_p(0)
    with torch.cuda.device(x.device):
        _native.check(lib.rfk_decode_u8(x.data_ptr(), out.data_ptr(), n, d, ld, sp, hp, code, _stream_ptr(stream)))
    return out


def resize_decode_u8(x, s_out, scale=None, shift=None, out_dtype=torch.uint8, stream=None, out=None):
    """Pillow-NEAREST resize (+ optional decode): ``uint8 [n, s, s(, ch)]`` -> ``[n, s_out*s_out*ch]``."""
    lib = _native.load()
    n, s_in = x.shape[0], x.shape[1]
    ch = x.shape[3] if x.dim() == 4 else 1
    d = s_out * s_out * ch
    if out is None:
        out = torch.empty((n, d), dtype=out_dtype, device=x.device)
    else:
        out_dtype = out.dtype
        if tuple(out.shape) != (n, d) or not out.is_contiguous():
            raise ValueError('out must be a contiguous [{}, {}] tensor'.format(n, d))
    code = {torch.uint8: -1, torch.float32: _native.RFK_F32, torch.bfloat16: _native.RFK_BF16}[out_dtype]
    sp = ctypes.c_void_p(scale.data_ptr()) if scale is not None else ctypes.c_void_p(0)
    hp = ctypes.c_void_p(shift.data_ptr()) if shift is not None else ctypes.c_void_p(0)
    with torch.cuda.device(x.device):
        _native.check(lib.rfk_resize_decode_u8(x.data_ptr(), out.data_ptr(), n, s_in, s_out, ch, d, sp, hp, code,
                                               _stream_ptr(stream)))
    return out


def frames_as_uint8(batch, frame_shape=None):
    """Binary ingress: ``uint8`` frames without a per-pixel Python conversion.

    ``batch`` is an ``ndarray`` of dtype uint8 shaped ``[n, S, S(, ch)]`` (or ``[n, d]``), or a
    bytes-like object holding ``n`` frames back to back (``frame_shape`` gives one frame's shape).
    Returns a C-contiguous ``uint8 [n, ...]`` array that shares the caller's memory when it can.
    """
    if isinstance(batch, (bytes, bytearray, memoryview)):
        if frame_shape is None:
            raise ValueError('a bytes batch needs frame_shape')
        per = int(np.prod(frame_shape))
        a = np.frombuffer(batch, dtype=np.uint8)
        if per == 0 or a.size % per:
            raise ValueError('{} bytes are not a whole number of {}-byte frames'.format(a.size, per))
        return a.reshape((a.size // per,) + tuple(int(v) for v in frame_shape))
    a = np.asarray(batch)
    if a.dtype != np.uint8:
        raise TypeError('binary batches must be uint8 (got {})'.format(a.dtype))
    return np.ascontiguousarray(a)


class HostPipeline:
    """Pinned-host hand-off around an :class:`Ensemble` (or :class:`Trial`): what replaces the Redis
    lists of rafiki/cache/cache.py:36-63 between the predictor and co-located workers.

    ``serve(batch) -> float32 numpy [n, C]`` copies the uint8 frames from pinned host memory to the
    device, resizes them on the GPU when the callers' frames are not the size the trials were
    trained on (``input_shape``; Pillow-NEAREST, ``rfk_resize_decode_u8``: the reference's
    ``resize_as_images``, rafiki/model/dataset.py:76-78), runs the engine, and copies the
    probabilities back, all on one stream; with ``depth`` > 1 slots, ``submit``/``collect`` keep
    several batches in flight so the copies of one batch overlap the kernels of another.  Every slot
    has its own stream, buffers AND kernel workspace, so slots never share scratch memory.
    """

    def __init__(self, engine, max_batch, depth=2, input_shape=None):
        _require_cuda()
        self.engine = engine
        self.device = engine.device
        self.max_batch = int(max_batch)
        self.depth = int(depth)
        d, c = engine.d, engine.n_classes
        self.input_shape = tuple(int(v) for v in input_shape) if input_shape is not None else None
        self._resize = None
        d_host = d
        if self.input_shape is not None:
            s_in = self.input_shape[0]
            ch = self.input_shape[2] if len(self.input_shape) == 3 else 1
            if len(self.input_shape) < 2 or self.input_shape[1] != s_in:
                raise ValueError('frames must be square: {}'.format(self.input_shape))
            if engine.image_size is None or ch != engine.channels:
                raise ValueError('frames {} do not match the trials ({} channels)'.format(self.input_shape, engine.channels))
            if s_in != engine.image_size:
                can_fold = getattr(engine, 'can_fold_resize', None)
                if can_fold is None:
                    can_fold = getattr(engine, 'compute', 'f32') not in ('f32', 'fp32')
                if can_fold:       # tensor-core trials: the resize folds into the first layer, frames go in as they are
                    engine = engine.for_input(s_in)
                    self.engine = engine
                    d = engine.d
                else:
                    self._resize = (s_in, engine.image_size, ch)
                d_host = s_in * s_in * ch
        self.d_host = d_host
        self._slots = []
        ws_bytes = max(engine.workspace_bytes(n) for n in {self.max_batch, min(self.max_batch, 128), 1})
        for _ in range(self.depth):
            self._slots.append({
                'h_in': torch.empty((self.max_batch, d_host), dtype=torch.uint8).pin_memory(),
                'h_out': torch.empty((self.max_batch, c), dtype=torch.float32).pin_memory(),
                'd_raw': (torch.empty((self.max_batch, d_host), dtype=torch.uint8, device=self.device)
                          if self._resize else None),
                'd_in': torch.empty((self.max_batch, d), dtype=torch.uint8, device=self.device),
                'd_out': torch.empty((self.max_batch, c), dtype=torch.float32, device=self.device),
                'stream': torch.cuda.Stream(device=self.device),
                'ws': Workspace(self.device, ws_bytes),
                'done': torch.cuda.Event(),
                'n': 0,
            })
        self._next = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def stage(self, slot_idx, batch):
        """Copy ``uint8`` frames (ndarray ``[n, ...]`` or bytes) into the slot's pinned input buffer."""
        s = self._slots[slot_idx]
        a = frames_as_uint8(batch, self.input_shape if self.input_shape is not None else (self.d_host,))
        n = a.shape[0]
        if n > self.max_batch:
            raise ValueError('batch of {} exceeds max_batch {}'.format(n, self.max_batch))
        if a.size != n * self.d_host:
            raise ValueError('frames have {} bytes each, the pipeline expects {}'.format(a.size // max(n, 1), self.d_host))
        s['h_in'][:n].copy_(torch.from_numpy(a.reshape(n, -1)))
        s['n'] = n

    def launch(self, slot_idx, n=None):
        """H2D + kernels + D2H of one slot, asynchronously on the slot's stream."""
        s = self._slots[slot_idx]
        n = s['n'] if n is None else n
        s['n'] = n
        with torch.cuda.stream(s['stream']):
            if self._resize:
                s_in, s_out, ch = self._resize
                s['d_raw'][:n].copy_(s['h_in'][:n], non_blocking=True)
                resize_decode_u8(s['d_raw'][:n].view((n, s_in, s_in) + ((ch,) if ch > 1 else ())), s_out,
                                 out=s['d_in'][:n], stream=s['stream'])
            else:
                s['d_in'][:n].copy_(s['h_in'][:n], non_blocking=True)
            self.engine.forward(s['d_in'][:n], out=s['d_out'][:n], stream=s['stream'], workspace=s['ws'])
            s['h_out'][:n].copy_(s['d_out'][:n], non_blocking=True)
            s['done'].record(s['stream'])
        self.h2d_bytes = n * self.d_host
        self.d2h_bytes = n * self.engine.n_classes * 4

    def wait(self, slot_idx):
        s = self._slots[slot_idx]
        s['done'].synchronize()
        return s['h_out'][:s['n']]

    def serve(self, batch):
        self.stage(0, batch)
        self.launch(0)
        return self.wait(0).numpy().copy()
