"""Host-side sharding of the search path across GPUs.

Records are independent, so N ranks take N contiguous ranges of the input; nothing is exchanged on
the data path.  The only cross-rank step is the final match count (one integer per rank, summed) and
the ordered concatenation of the per-rank outputs.  Pure host logic: no GPU, no matcher in here.
"""
import ctypes as C


def record_ranges(n_records, world):
    """Contiguous record ranges [first, first + count) per rank; sizes differ by at most one."""
    base, extra = divmod(n_records, world)
    out, first = [], 0
    for r in range(world):
        cnt = base + (1 if r < extra else 0)
        out.append((first, cnt))
        first += cnt
    return out


def byte_shards(buf, world, find_cut):
    """Cuts a raw FASTQ buffer into `world` contiguous byte ranges that end on record boundaries.
    `find_cut(view, n)` is fxg_find_record_cut (largest record-aligned prefix length <= n).
    Returns [(offset, length)] covering the buffer exactly, in order; trailing shards may be empty."""
    n = len(buf)
    cuts = [0]
    for r in range(1, world):
        target = n * r // world
        if target <= cuts[-1]:
            cuts.append(cuts[-1])
            continue
        c = find_cut(buf, target)
        cuts.append(max(c, cuts[-1]))
    cuts.append(n)
    return [(cuts[i], cuts[i + 1] - cuts[i]) for i in range(world)]


def reduce_counts(local_count, dist=None):
    """Sum of the per-rank match counts (the path's only collective)."""
    if dist is None or not dist.is_initialized():
        return int(local_count)
    import torch
    t = torch.tensor([int(local_count)], dtype=torch.int64)
    backend = dist.get_backend()
    if backend == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return int(t.item())


def gather_in_rank_order(local_bytes, dist=None):
    """Per-rank outputs concatenated in rank order == input order (rank 0 gets the whole; others b"")."""
    if dist is None or not dist.is_initialized():
        return local_bytes
    world, rank = dist.get_world_size(), dist.get_rank()
    parts = [None] * world if rank == 0 else None
    dist.gather_object(local_bytes, parts, dst=0)
    return b"".join(parts) if rank == 0 else b""
