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

    Every rank owns two partial-mean buffers (alternating by step), two result buffers and one
    flag block; the buffers are exported with CUDA IPC (through ``torch.multiprocessing``'s tensor
    reduction) and opened by every other rank, so a kernel on GPU r can load rank j's partials and
    store into the root's result directly.  ``reduce(step_buffer)`` enqueues, on the current stream:
    the reduce-scatter-plus-gather kernel on every rank and, on the root, the wait for all slices.
    """

    def __init__(self, n, n_classes, device, root=0, group=None, grid_blocks=32):
        from torch.multiprocessing.reductions import reduce_tensor
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
        self.partial = [torch.zeros((n, n_classes), dtype=torch.float32, device=self.device) for _ in range(2)]
        self.result = [torch.zeros((n, n_classes), dtype=torch.float32, device=self.device) for _ in range(2)]
        self.flags = torch.zeros(2 * self.world + 1, dtype=torch.int32, device=self.device)
        torch.cuda.synchronize(self.device)
        mine = [reduce_tensor(t) for t in self.partial + self.result + [self.flags]]
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=group)
        # keep the rebuilt tensors alive: they hold the IPC mappings
        self._peer = []
        for r, handles in enumerate(everyone):
            if r == self.rank:
                self._peer.append(self.partial + self.result + [self.flags])
            else:
                self._peer.append([fn(*args) for (fn, args) in handles])
        vp = ctypes.c_void_p
        self._probs_ptrs = [(vp * self.world)(*[self._peer[r][k].data_ptr() for r in range(self.world)]) for k in (0, 1)]
        self._flag_ptrs = (vp * self.world)(*[self._peer[r][4].data_ptr() for r in range(self.world)])
        self._root_out = [self._peer[self.root][2 + k].data_ptr() for k in (0, 1)]
        dist.barrier(group=group)

    def next_partial(self):
        """The buffer this rank's fused kernel must write for the coming step."""
        return self.partial[(self.step + 1) & 1]

    def reduce(self):
        """Enqueue the cross-rank mean of the buffers handed out by ``next_partial``; returns the
        result tensor (meaningful on the root, once the stream has passed this point)."""
        self.step += 1
        k = self.step & 1
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            self._check(self._lib.rfk_peer_reduce(self._probs_ptrs[k], self._flag_ptrs, self.world, self.rank, self.root,
                                                  self.step, self.count, self._root_out[k], self.grid_blocks, stream))
            if self.rank == self.root:
                self._check(self._lib.rfk_peer_wait(self.flags.data_ptr(), self.world, self.step, stream))
        return self.result[k]


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
