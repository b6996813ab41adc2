"""Frame-level sharding across the GPUs of one box: independent frames / stereo
pairs are split into contiguous blocks, one block per rank; there is no data-path collective.
The only communication is the host-side gather of per-frame outputs (torch.distributed
all_gather_object when several processes are used)."""


def shard_range(n_items, rank, world_size):
    """Contiguous block [lo, hi) of rank `rank`; the first n_items % world_size ranks get one extra."""
    if world_size < 1 or not (0 <= rank < world_size) or n_items < 0:
        raise ValueError("bad shard arguments")
    base, extra = divmod(n_items, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_per_frame(local_results, n_items, dist=None):
    """Host gather: every rank passes the list of results for its shard_range; returns the full
    list in frame order on every rank.  With dist=None (single process) returns the input."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        assert len(local_results) == n_items
        return list(local_results)
    world = dist.get_world_size()
    parts = [None] * world
    dist.all_gather_object(parts, list(local_results))
    out = []
    for r in range(world):
        lo, hi = shard_range(n_items, r, world)
        assert len(parts[r]) == hi - lo, "rank %d returned %d results for [%d, %d)" % (r, len(parts[r]), lo, hi)
        out.extend(parts[r])
    return out
