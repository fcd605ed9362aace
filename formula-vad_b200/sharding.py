"""Stream sharding across GPUs: streams are independent pipelines -- the reference runs
them as independent OS threads (src/simulator.zig:221-232) -- so they partition statically over ranks,
one process per GPU, with NO collective on the data path.  The only exchange is the "host gather" of the
per-stream SpeechSegment lists (tens of bytes each) at the end, done here with torch.distributed
object gathering (gloo on CPU, NCCL-backed process groups on the GPU box)."""
from __future__ import annotations


def partition_streams(n_streams: int, world_size: int, rank: int) -> range:
    """Contiguous block partition; the first (n_streams % world_size) ranks take one extra stream."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad rank/world_size")
    base, extra = divmod(n_streams, world_size)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def gather_segments(local: dict, dst: int = 0, group=None):
    """local: {global_stream_id: [segment tuples]} of this rank.  Returns the merged dict on rank
    `dst` (None elsewhere).  Without an initialised process group it is the identity."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return dict(local)
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    out = [None] * world if rank == dst else None
    dist.gather_object(local, out, dst=dst, group=group)
    if rank != dst:
        return None
    merged = {}
    for part in out:
        for k, v in part.items():
            if k in merged:
                raise ValueError(f"stream {k} reported by two ranks")
            merged[k] = v
    return merged
