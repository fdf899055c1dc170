"""Data-parallel plumbing for the quantiser: one process per GPU, replicated codebook,
row-sharded latents, ONE collective per step -- all_reduce(SUM) over the contiguous
[code_sum (K*D) | code_cnt (K)] fp32 buffer -- after which every rank applies the identical EMA update.
The reference has no distributed code (single GPU, frl/config/frl_training_v1.yaml:26-27); this is the
shard point north_star prescribes.  Backend-agnostic (NCCL on the GPUs, gloo in the CPU tests)."""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced row shard [a, b) of rank `rank` (first n_rows % world ranks get one more)."""
    base, rem = divmod(int(n_rows), int(world))
    a = rank * base + min(rank, rem)
    return a, a + base + (1 if rank < rem else 0)


def dist_active(group=None) -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def allreduce_code_stats(stats: torch.Tensor, group=None) -> torch.Tensor:
    """In-place SUM over ranks of the packed statistics buffer; a no-op for a single process."""
    if dist_active(group):
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def codebook_checksum(codebook: torch.Tensor) -> torch.Tensor:
    """Order-independent fingerprint used to assert that replicas have not drifted apart."""
    bits = codebook.detach().contiguous().view(torch.int32).to(torch.int64)
    return torch.stack([bits.sum(), (bits * bits % 1000003).sum()])


def assert_replicas_identical(codebook: torch.Tensor, group=None) -> None:
    if not dist_active(group):
        return
    mine = codebook_checksum(codebook)
    world = dist.get_world_size(group)
    allc = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allc, mine, group=group)
    for r, c in enumerate(allc):
        if not torch.equal(c, allc[0]):
            raise RuntimeError(f"codebook replica of rank {r} differs from rank 0")
