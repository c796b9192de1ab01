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

    def __init__(self, n, n_classes, device, root=0, group=None, grid_blocks=148, overlap=False):
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
        # the reduce runs on its own stream so that it overlaps the next batch's fused kernel
        self.overlap = bool(overlap)
        self._side = torch.cuda.Stream(device=self.device) if self.overlap else None
        self._ev_ready = [torch.cuda.Event() for _ in range(2)]
        self._ev_done = [torch.cuda.Event() for _ in range(2)]
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

    def _as_tensor(self, ptr, shape):
        """A float32 torch view of raw device memory owned by this object."""
        class _Holder:
            pass
        hld = _Holder()
        hld.__cuda_array_interface__ = {'shape': tuple(shape), 'typestr': '<f4', 'data': (int(ptr), False), 'version': 2}
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
        k = (self.step + 1) & 1
        if self.overlap and self.step >= 2:
            torch.cuda.current_stream(self.device).wait_event(self._ev_done[k])
        return self.partial[k]

    def reduce(self):
        """Enqueue the cross-rank mean of the buffers handed out by ``next_partial``; returns the
        result tensor (meaningful on the root, once the stream has passed this point)."""
        self.step += 1
        k = self.step & 1
        cur = torch.cuda.current_stream(self.device)
        run_on = self._side if self.overlap else cur
        if self.overlap:
            self._ev_ready[k].record(cur)               # the partials of this step are written
            run_on.wait_event(self._ev_ready[k])
        stream = ctypes.c_void_p(run_on.cuda_stream)
        if torch.cuda.current_device() != self.device.index:
            torch.cuda.set_device(self.device)       # one process per GPU: this rank only ever uses its own device
        self._check(self._lib.rfk_peer_reduce(self._probs_ptrs[k], self._flag_ptrs, self.world, self.rank, self.root,
                                              self.step, self.count, self._root_out[k], self.grid_blocks, stream))
        if self.rank == self.root:
            self._check(self._lib.rfk_peer_wait(self._flags_ptr, self.world, self.step, stream))
        if self.overlap:
            self._ev_done[k].record(run_on)
        return self.result[k]

    def wait(self, k=None):
        """Make the current stream wait for the reduce of the last step (or of parity k)."""
        if self.overlap:
            torch.cuda.current_stream(self.device).wait_event(self._ev_done[self.step & 1 if k is None else k])


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
        self.overlap = False
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
