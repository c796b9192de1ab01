"""Multi-GPU serving of one ensemble: trials placed on ranks, partial means summed onto the
predictor rank.

This mirrors the reference's deployment -- one inference worker service per best trial
(rafiki/admin/services_manager.py:62-67), every worker sees every query, the predictor gathers
one answer per worker and averages (rafiki/predictor/predictor.py:35-70) -- with the Redis
round trips replaced by one collective: rank r runs its trials' fused kernel with
``total_trials = W`` (so its output is the partial mean ``sum_{t on r} p_t / W``) and the
partials are summed onto ``root`` with ``torch.distributed.reduce``.  One process per GPU; the
process group (NCCL on GPUs, gloo in the CPU tests) is the caller's.
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


class PeerReducer:
    """Cross-rank mean over NVLink peer memory (``rfk_peer_reduce``): the collective fused into
    kernel 3 instead of an NCCL reduce.

    Every rank owns one device block (``rfk_peer_alloc``) holding two partial-mean buffers
    (alternating by step), two result buffers and a flag block; its CUDA IPC handle goes to every
    other rank, which maps it with its own device current (``rfk_peer_open``), so a kernel on GPU r
    can load rank j's partials and store into the root's result directly.  ``reduce(step_buffer)`` enqueues, on the current stream:
    the reduce-scatter-plus-gather kernel on every rank and, on the root, the wait for all slices.
    """

    def __init__(self, n, n_classes, device, root=0, group=None, grid_blocks=148):
        from . import _native
        self._lib = _native.load()
        self._check = _native.check
        self.device = torch.device(device)
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.root = root
        self.count = int(n) * int(n_classes)
        if self.count % 4:
            raise ValueError('n * n_classes must be a multiple of 4 for the peer reduce')
        self.grid_blocks = int(grid_blocks)
        self.step = 0
        # Everything of one rank runs on ONE stream: the fused kernel of step s+2 is then ordered
        # after this rank's reduce of step s+1, which waited for every peer's "ready" of step s+1,
        # which each peer raised after its own reduce of step s -- so nobody still reads the
        # partials of step s when they are overwritten.  (A side-stream variant broke that chain.)
        # one block per rank: partial[0] | partial[1] | result[0] | result[1] | flags
        buf_bytes = (self.count * 4 + 255) // 256 * 256
        self._offsets = [0, buf_bytes, 2 * buf_bytes, 3 * buf_bytes, 4 * buf_bytes]
        total = 4 * buf_bytes + 256
        base = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(self.device):
            self._check(self._lib.rfk_peer_alloc(total, ctypes.byref(base), handle))
        self._base = base.value
        everyone = [None] * self.world
        dist.all_gather_object(everyone, (bytes(handle.raw), self.device.index), group=group)
        self._bases, self._opened = [], []
        with torch.cuda.device(self.device):      # map with MY device current: my kernels will touch it
            for r, (h, peer_dev) in enumerate(everyone):
                if r == self.rank:
                    self._bases.append(self._base)
                    continue
                self._check(self._lib.rfk_peer_enable(int(peer_dev)))
                p = ctypes.c_void_p()
                self._check(self._lib.rfk_peer_open(ctypes.create_string_buffer(h, 64), ctypes.byref(p)))
                self._bases.append(p.value)
                self._opened.append(p.value)
        vp = ctypes.c_void_p
        self._probs_ptrs = [(vp * self.world)(*[b + self._offsets[k] for b in self._bases]) for k in (0, 1)]
        self._flag_ptrs = (vp * self.world)(*[b + self._offsets[4] for b in self._bases])
        self._root_out = [self._bases[self.root] + self._offsets[2 + k] for k in (0, 1)]
        self.partial = [self._as_tensor(self._base + self._offsets[k], (n, n_classes)) for k in (0, 1)]
        self.result = [self._as_tensor(self._base + self._offsets[2 + k], (n, n_classes)) for k in (0, 1)]
        self._flags_ptr = self._base + self._offsets[4]
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
            for p in self._opened:
                self._lib.rfk_peer_close(p)
            self._lib.rfk_peer_free(self._base)
        self._base = None

    def next_partial(self):
        """The buffer this rank's fused kernel must write for the coming step.  (Its previous
        contents, two steps old, have been consumed: the current stream waits for that reduce.)"""
        return self.partial[(self.step + 1) & 1]

    def reduce(self):
        """Enqueue the cross-rank mean of the buffers handed out by ``next_partial``; returns the
        result tensor (meaningful on the root, once the stream has passed this point)."""
        self.step += 1
        k = self.step & 1
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        if torch.cuda.current_device() != self.device.index:
            torch.cuda.set_device(self.device)       # one process per GPU: this rank only ever uses its own device
        self._check(self._lib.rfk_peer_reduce(self._probs_ptrs[k], self._flag_ptrs, self.world, self.rank, self.root,
                                              self.step, self.count, self._root_out[k], self.grid_blocks, stream))
        if self.rank == self.root:
            self._check(self._lib.rfk_peer_wait(self._flags_ptr, self.world, self.step, stream))
        return self.result[k]

    def wait(self, k=None):
        """Kept for callers of the earlier interface: the reduce runs on the calling stream."""


class PeerGather(PeerReducer):
    """The same collective fused into the trial kernel itself: a rank's fused kernel is handed an
    output pointer that lies in the ROOT's memory (its slot of the root's block), so the epilogue
    stores its partial means across NVLink tile by tile while the tensor cores work on the next tile.
    After the kernel each rank raises a flag on the root; the root waits for all flags, adds the G
    slots (``rfk_ensemble_sum``, float64) and acknowledges, which frees the slots for the step after
    next.  No separate communication kernel runs on the non-root ranks at all.
    """

    def __init__(self, n, n_classes, device, root=0, group=None):
        from . import _native
        self._lib = _native.load()
        self._check = _native.check
        self.device = torch.device(device)
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.root = root
        self.n, self.n_classes = int(n), int(n_classes)
        self.count = self.n * self.n_classes
        self.step = 0
        G = self.world
        buf_bytes = (self.count * 4 + 255) // 256 * 256
        self._buf_bytes = buf_bytes
        self._flags_off = (2 * G + 2) * buf_bytes           # [0,G): arrived (on the root); [G]: consumed (on each rank)
        total = self._flags_off + 256
        base = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(self.device):
            self._check(self._lib.rfk_peer_alloc(total, ctypes.byref(base), handle))
        self._base = base.value
        everyone = [None] * G
        dist.all_gather_object(everyone, (bytes(handle.raw), self.device.index), group=group)
        self._bases, self._opened = [], []
        with torch.cuda.device(self.device):
            for r, (h, peer_dev) in enumerate(everyone):
                if r == self.rank:
                    self._bases.append(self._base)
                    continue
                self._check(self._lib.rfk_peer_enable(int(peer_dev)))
                p = ctypes.c_void_p()
                self._check(self._lib.rfk_peer_open(ctypes.create_string_buffer(h, 64), ctypes.byref(p)))
                self._bases.append(p.value)
                self._opened.append(p.value)
        root_base = self._bases[self.root]
        vp = ctypes.c_void_p
        # my slot in the root's block, per parity, as a tensor my fused kernel can be pointed at
        self.partial = [self._as_tensor(root_base + (k * G + self.rank) * buf_bytes, (n, n_classes)) for k in (0, 1)]
        self.result = [self._as_tensor(self._base + (2 * G + k) * buf_bytes, (n, n_classes)) for k in (0, 1)]
        self._my_arrived = (vp * 1)(root_base + self._flags_off + 4 * self.rank)
        self._consumed_local = self._base + self._flags_off + 4 * G
        if self.rank == self.root:
            self._slot_ptrs = [(vp * G)(*[self._base + (k * G + r) * buf_bytes for r in range(G)]) for k in (0, 1)]
            self._ack_ptrs = (vp * G)(*[b + self._flags_off + 4 * G for b in self._bases])
        dist.barrier(group=group)

    def next_partial(self):
        k = (self.step + 1) & 1
        if self.step >= 2:      # the root must have added the slot's previous contents (two steps ago)
            stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            with torch.cuda.device(self.device):
                self._check(self._lib.rfk_peer_wait_flags(self._consumed_local, 1, self.step - 1, stream))
        return self.partial[k]

    def reduce(self):
        self.step += 1
        k = self.step & 1
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        G = self.world
        with torch.cuda.device(self.device):
            self._check(self._lib.rfk_peer_signal(self._my_arrived, 1, self.step, stream))
            if self.rank == self.root:
                self._check(self._lib.rfk_peer_wait_flags(self._base + self._flags_off, G, self.step, stream))
                self._check(self._lib.rfk_ensemble_sum(self._slot_ptrs[k], G, self.n, self.n_classes,
                                                      self.result[k].data_ptr(), stream))
                self._check(self._lib.rfk_peer_signal(self._ack_ptrs, G, self.step, stream))
        return self.result[k]

    def wait(self, k=None):
        pass


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

    _as_tensor = PeerReducer._as_tensor

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

    def forward_peer(self, x, reducer):
        """Local kernel into the reducer's partial buffer, then the fused peer-memory reduce; all
        on the current stream, no host synchronisation.  Returns the root's result tensor."""
        part = reducer.next_partial()
        if self.local is not None:
            self.local.forward(x, out=part)
        else:
            part.zero_()
        return reducer.reduce()

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
