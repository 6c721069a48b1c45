"""Host-side helpers for running the path on several GPUs of one box (one process per GPU).

Queries are independent and the index is read-only, so kNN / radius search shard by
contiguous query blocks with no collective.  ICP shards the source cloud and
needs one small all-reduce per iteration: the partial sums of the normal equations
(registration.cpp:788-818 builds them from ALL pairs; here every rank contributes its pairs).
"""

ICP_SUM_SLOTS = 31  # doubles [0, 31) are summed; slot 31 carries an int64 that takes the MAX


def shard_range(n, world_size, rank):
    """Contiguous block [lo, hi) of ``n`` items owned by ``rank``; blocks tile [0, n) exactly and
    differ in size by at most one item."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError("bad rank %r / world size %r" % (rank, world_size))
    return rank * n // world_size, (rank + 1) * n // world_size


def allreduce_icp_sums(sums, dist, group=None):
    """All-reduce the PCNN_ICP_NSUMS doubles of ``pcnn_icp_sums_dev`` in place.

    ``sums``: float64 tensor of 32 elements (CUDA with NCCL, CPU with gloo).  Elements [0, 31)
    are partial sums (A^T A, A^T b, b^T b, pair count / centroids and cross terms); element 31
    holds, bit-cast to int64, (global index of the last kept pair << 32 | float bits of its d2)
    and takes the MAX, which reproduces the reference's "loss of the last pair" (:939).
    """
    import torch
    if sums.dtype != torch.float64 or sums.numel() != 32:
        raise ValueError("expected 32 float64 sums")
    dist.all_reduce(sums[:ICP_SUM_SLOTS], op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(sums[ICP_SUM_SLOTS:].view(torch.int64), op=dist.ReduceOp.MAX, group=group)
    return sums
