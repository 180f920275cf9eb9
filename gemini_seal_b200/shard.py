"""Ciphertext sharding across ranks: ciphertexts are independent, so a batch is cut into
contiguous shards, one per GPU / process, and nothing is exchanged on the data path.  The only
collectives are the plumbing around a job: scatter the inputs from rank 0, gather the results (or
their checksums) back.  Backend-agnostic (`nccl` on the GPU box, `gloo` in the CPU tests)."""
import torch
import torch.distributed as dist


def shard_bounds(batch, world, rank):
    """Contiguous, balanced [lo, hi) of `batch` units for `rank` (first batch % world ranks get one more)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    base, extra = divmod(batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def scatter_batch(full, per_rank_shape, dtype, device, src=0):
    """Rank `src` holds `full` = [world * B, ...]; every rank receives its [B, ...] shard."""
    world, rank = dist.get_world_size(), dist.get_rank()
    out = torch.empty(per_rank_shape, dtype=dtype, device=device)
    if world == 1:
        out.copy_(full)
        return out
    chunks = [c.contiguous() for c in full.chunk(world)] if rank == src else None
    dist.scatter(out, chunks, src=src)
    return out


def gather_batch(shard, dst=0):
    """Inverse of scatter_batch: rank `dst` gets the concatenation of all shards, others get None."""
    world, rank = dist.get_world_size(), dist.get_rank()
    if world == 1:
        return shard
    bufs = [torch.empty_like(shard) for _ in range(world)] if rank == dst else None
    dist.gather(shard.contiguous(), bufs, dst=dst)
    return torch.cat(bufs) if rank == dst else None
