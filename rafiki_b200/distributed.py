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
import torch
import torch.distributed as dist


def trials_of_rank(n_trials, world_size, rank):
    """Round-robin placement: trial t lives on rank t % world_size (one trial per GPU when
    n_trials == world_size; ceil(W / G) per GPU otherwise)."""
    if not (0 <= rank < world_size):
        raise ValueError('rank {} outside world of {}'.format(rank, world_size))
    return [t for t in range(n_trials) if t % world_size == rank]


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
