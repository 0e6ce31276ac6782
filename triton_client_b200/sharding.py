"""Sharding of the marshalling path across the GPUs of one box.

The path has no exchange step: images are independent and BYTES serialisation is concatenative
(``ser(A || B) == ser(A) || ser(B)``). Every rank therefore works on a contiguous range of units
and fills its own registered region; the only cross-rank quantity is, optionally, each rank's base
offset inside one logical BYTES tensor, an exclusive sum of R integers. No collective touches
the data path; ``torch.distributed`` is used only to agree on those R integers (and, in bench.py,
for the barrier around the timed region).
"""


def shard_range(n_units, rank, world_size):
    """Contiguous, balanced range [lo, hi) of `n_units` for `rank`: the first n % world ranks get
    one extra unit (ranges are ordered by rank, so concatenating the per-rank outputs in rank order
    reproduces the unsharded result)."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad rank %r / world size %r" % (rank, world_size))
    base, extra = divmod(int(n_units), int(world_size))
    lo = rank * base + min(rank, extra)
    hi = lo + base + (1 if rank < extra else 0)
    return lo, hi


def exclusive_offsets(sizes):
    """Base offset of every rank's piece inside the concatenated tensor."""
    out, run = [], 0
    for s in sizes:
        out.append(run)
        run += int(s)
    return out, run


def gather_sizes(local_size, group=None):
    """All ranks' `local_size` (python ints, rank order) through torch.distributed.all_gather on the
    current backend (gloo on CPU, nccl on GPU); a single-process run returns [local_size]."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return [int(local_size)]
    world = dist.get_world_size(group)
    device = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    mine = torch.tensor([int(local_size)], dtype=torch.int64, device=device)
    allv = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allv, mine, group=group)
    return [int(t.item()) for t in allv]


def bytes_base_offset(local_serialized_bytes, group=None):
    """(my base offset, total bytes) of this rank's serialized piece inside the logical tensor."""
    import torch.distributed as dist

    sizes = gather_sizes(local_serialized_bytes, group)
    offs, total = exclusive_offsets(sizes)
    rank = dist.get_rank(group) if (dist.is_available() and dist.is_initialized()) else 0
    return offs[rank], total
