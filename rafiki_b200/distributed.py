"""Multi-GPU serving of one ensemble: trials placed on ranks, partial means summed onto the
predictor rank.

This mirrors the reference's deployment -- one inference worker service per best trial
(rafiki/admin/services_manager.py:62-67), every worker sees every query, the predictor gathers
one answer per worker and averages (rafiki/predictor/predictor.py:35-70) -- with the Redis
round trips replaced by one collective: rank r runs its trials' fused kernel with
``total_trials = W`` (so its output is the partial mean ``sum_{t on r} p_t / W``) and the
partials are summed onto ``root``.  On GPUs the collective is this library's own
(:class:`PeerEnsemble`: partial means stored straight into peer memory over NVLink from the trial
kernel's epilogue, summed by their owners, ``rfk_peer_ensemble_step*``); ``torch.distributed`` is
used for the rendezvous only (exchange of the CUDA-IPC handles, barriers).
:class:`DistributedEnsemble` is the same composition over ``torch.distributed.reduce`` (gloo in
the CPU tests).  One process per GPU; the process group is the caller's.
"""
import ctypes

import torch
import torch.distributed as dist


def trials_of_rank(n_trials, world_size, rank):
    """Round-robin placement: trial t lives on rank t % world_size (one trial per GPU when
    n_trials == world_size; ceil(W / G) per GPU otherwise)."""
    if not (0 <= rank < world_size):
        raise ValueError('rank {} outside world of {}'.format(rank, world_size))
    return [t for t in range(n_trials) if t % world_size == rank]


def slice_rows(n, world_size, rank):
    """Rows ``[lo, hi)`` of an ``n``-row batch that ``rank`` uploads (``rfk_peer_slice_rows``: slices
    start at multiples of 16 rows so that every slice is 16-byte aligned for any row length)."""
    if not (0 <= rank < world_size):
        raise ValueError('rank {} outside world of {}'.format(rank, world_size))
    units = (n + 15) // 16
    return min(units * rank // world_size * 16, n), min(units * (rank + 1) // world_size * 16, n)


class PeerEnsemble:
    """The trial-per-GPU ensemble with ONE native call per step (``rfk_peer_ensemble_step``).

    Mirrors one ``Predictor.predict`` round (rafiki/predictor/predictor.py:35-70) across ranks:
    every rank holds some trials and sees every query; the mean over all trials lands on ``root``.
    ``local`` is this rank's :class:`rafiki_b200.engine.Ensemble` (built with ``total_trials`` = all
    trials) or ``None`` for a rank without a trial.

    * ``step(x)``: fused kernel on the device-resident batch ``x`` -> peer-memory reduce -> (root)
      wait; returns the root's result tensor.
    * ``upload(batch_u8_pinned)`` + ``gather()``: each rank copies only ITS rows of the batch
      from pinned host memory (its own PCIe link) and the other slices arrive over NVLink
      (``rfk_peer_allgather_u8``); returns the whole batch on this device, ready for ``step``.
    Everything is enqueued on the current stream; the host never synchronises.
    """

    def __init__(self, local, n_trials, max_n, d, n_classes, device, root=0, group=None):
        from . import _native
        self._lib = _native.load()
        self._check = _native.check
        self.local = local
        self.n_trials = int(n_trials)
        self.device = torch.device(device)
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.root = root
        self.max_n, self.d, self.n_classes = int(max_n), int(d), int(n_classes)
        nbytes = ctypes.c_size_t()
        self._check(self._lib.rfk_peer_block_bytes(self.world, self.max_n, self.d, self.n_classes, ctypes.byref(nbytes)))
        base = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(self.device):
            self._check(self._lib.rfk_peer_alloc(nbytes.value, ctypes.byref(base), handle))
        self._base = base.value
        everyone = [None] * self.world
        dist.all_gather_object(everyone, (bytes(handle.raw), self.device.index), group=group)
        bases, self._opened = [], []
        with torch.cuda.device(self.device):      # map with MY device current: my kernels will touch it
            for r, (h, peer_dev) in enumerate(everyone):
                if r == self.rank:
                    bases.append(self._base)
                    continue
                self._check(self._lib.rfk_peer_enable(int(peer_dev)))
                p = ctypes.c_void_p()
                self._check(self._lib.rfk_peer_open(ctypes.create_string_buffer(h, 64), ctypes.byref(p)))
                bases.append(p.value)
                self._opened.append(p.value)
            ctx = ctypes.c_void_p()
            arr = (ctypes.c_void_p * self.world)(*bases)
            self._check(self._lib.rfk_peer_create(ctypes.byref(ctx), self.world, self.rank, self.root, arr,
                                                  self.max_n, self.d, self.n_classes))
        self._ctx = ctx.value
        self._handles = local._handles if local is not None else None
        self._w = len(local.trials) if local is not None else 0
        self._ws = None
        if local is not None:
            from .engine import Workspace
            self._ws = Workspace(self.device)
        self._plans = {}
        self._views = {}
        self._last_n = None
        self._res = ctypes.c_void_p()
        dist.barrier(group=group)

    def _as_tensor(self, ptr, shape, typestr='<f4', dtype=None):
        """A torch view (float32 by default) of raw device memory owned by this object."""
        class _Holder:
            pass
        hld = _Holder()
        hld.__cuda_array_interface__ = {'shape': tuple(shape), 'typestr': typestr, 'data': (int(ptr), False), 'version': 2}
        hld.owner = self
        return torch.as_tensor(hld, device=self.device)

    def close(self):
        if getattr(self, '_base', None) is None:
            return
        torch.cuda.synchronize(self.device)
        with torch.cuda.device(self.device):
            self._lib.rfk_peer_destroy(self._ctx)
            for p in self._opened:
                self._lib.rfk_peer_close(p)
            self._lib.rfk_peer_free(self._base)
        self._base = None

    def upload(self, host_batch):
        """Enqueue the H2D copy of this rank's rows of ``host_batch`` (pinned ``uint8 [n, d]`` tensor)
        into the slot the next ``gather`` reads.  Returns the number of bytes copied."""
        n = host_batch.shape[0]
        dst, lo, hi = ctypes.c_void_p(), ctypes.c_int64(), ctypes.c_int64()
        self._check(self._lib.rfk_peer_upload_slot(self._ctx, n, ctypes.byref(dst), ctypes.byref(lo), ctypes.byref(hi)))
        rows = hi.value - lo.value
        if rows > 0:
            view = self._as_tensor(dst.value, (rows, self.d), '|u1', torch.uint8)
            view.copy_(host_batch[lo.value:hi.value].reshape(rows, self.d), non_blocking=True)
        return rows * self.d

    def gather(self, n):
        """All-gather of the uploaded slices; returns the whole ``uint8 [n, d]`` batch on this device."""
        out = ctypes.c_void_p()
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        self._check(self._lib.rfk_peer_allgather_u8(self._ctx, n, ctypes.byref(out), stream))
        return self._as_tensor(out.value, (n, self.d), '|u1', torch.uint8)

    def step_pipelined(self, x):
        """Back-to-back batches (``rfk_peer_ensemble_step_pipelined``): enqueues this batch and returns the
        result tensor of the PREVIOUS call's batch (``None`` after the first call); with one trial per GPU
        the cross-rank sum of a batch rides inside the next batch's trial kernel.  ``flush()`` completes
        and returns the last one."""
        n = x.shape[0]
        plan = self._plans.get(n)
        if plan is None:
            ws, ws_bytes = (ctypes.c_void_p(0), 0)
            if self.local is not None:
                ws, ws_bytes = self._ws.get(self.local.workspace_bytes(n))
            plan = self._plans[n] = [ws, ws_bytes, {}]
        ws, ws_bytes, views = plan
        res = self._res
        rc = self._lib.rfk_peer_ensemble_step_pipelined(
            self._ctx, self._handles, self._w, self.n_trials, x.data_ptr(), n, ws, ws_bytes, ctypes.byref(res),
            ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        if rc != 0:
            self._check(rc)
        prev_n, self._last_n = self._last_n, n
        return self._view(res.value, prev_n)

    def set_inputs_resident(self, on=True):
        """Promise that the batches given to ``step_pipelined`` are complete before the PREVIOUS call was
        enqueued (resident batches): the trial kernel of a step may then start while the previous step's
        kernel is finishing (``rfk_peer_set_inputs_resident``).  Off by default."""
        self._check(self._lib.rfk_peer_set_inputs_resident(self._ctx, 1 if on else 0))

    def flush(self):
        res = ctypes.c_void_p()
        self._check(self._lib.rfk_peer_flush(self._ctx, ctypes.byref(res),
                                             ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        n, self._last_n = self._last_n, None
        return self._view(res.value, n)

    def _view(self, ptr, n):
        if not ptr or n is None:
            return None
        out = self._views.get((ptr, n))
        if out is None:
            out = self._views[(ptr, n)] = self._as_tensor(ptr, (n, self.n_classes))
        return out

    def step(self, x):
        """One serving step on the current stream.  The host side is kept to one native call: scratch
        sizes and the views of the two result buffers are cached per batch size (at eight ranks the
        kernels of a step take ~35 us, so every Python object built per step shows)."""
        n = x.shape[0]
        plan = self._plans.get(n)
        if plan is None:
            ws, ws_bytes = (ctypes.c_void_p(0), 0)
            if self.local is not None:
                ws, ws_bytes = self._ws.get(self.local.workspace_bytes(n))
            plan = self._plans[n] = [ws, ws_bytes, {}]
        ws, ws_bytes, views = plan
        res = self._res
        rc = self._lib.rfk_peer_ensemble_step(self._ctx, self._handles, self._w, self.n_trials, x.data_ptr(), n,
                                              ws, ws_bytes, ctypes.byref(res),
                                              ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        if rc != 0:
            self._check(rc)
        out = views.get(res.value)
        if out is None:
            out = views[res.value] = self._as_tensor(res.value, (n, self.n_classes))
        return out


class DistributedEnsemble:
    """``forward(x)`` on every rank -> the ensemble mean on ``root`` (partial sums elsewhere).

    ``local`` is this rank's engine over its own trials, already built with
    ``total_trials = n_trials`` (:class:`rafiki_b200.engine.Ensemble`), or ``None`` for a rank
    that holds no trial (W < G); such a rank contributes zeros.
    """

    def __init__(self, local, n_trials, n_classes, device, root=0, group=None):
        self.local = local
        self.n_trials = int(n_trials)
        self.n_classes = int(n_classes)
        self.device = torch.device(device)
        self.root = root
        self.group = group
        self.world_size = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def forward_async(self, x, out):
        """Local kernel on the current stream, then the reduce as an asynchronous collective (NCCL
        runs it on its own stream, ordered after the kernel).  Returns the work handle (``None``
        for a single rank); ``out`` must not be reused before ``handle.wait()``.  With two ``out``
        buffers used in turn the reduce of batch i overlaps the kernel of batch i+1."""
        if self.local is not None:
            self.local.forward(x, out=out)
        else:
            out.zero_()
        if self.world_size > 1:
            return dist.reduce(out, dst=self.root, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return None

    def forward(self, x, out=None):
        n = x.shape[0]
        if out is None:
            out = torch.empty((n, self.n_classes), dtype=torch.float32, device=self.device)
        if self.local is not None:
            self.local.forward(x, out=out)
        else:
            out.zero_()
        if self.world_size > 1:
            dist.reduce(out, dst=self.root, op=dist.ReduceOp.SUM, group=self.group)
        return out
