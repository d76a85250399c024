"""One input over several GPUs: contiguous parts, one rank (process + GPU) each, stitched on the host.

north_star / SURVEY.md §8e: shards are independent windows, so a rank needs nothing from the others
but the two bytes before its part (literal context).  There is no data-path collective: every rank
encodes its part with `Context.encode_part` (bgpu_encode_part) and the compressed parts are gathered
to one rank over the host (torch.distributed object gather; gloo on CPU boxes, the default group
otherwise) and concatenated — parts end byte-aligned (empty metadata block), the first one carries
the window header and the last one ISLAST, so the concatenation is one valid stream.

The reference is single-process (src/encode/encode.ts:181-281 loops over metablocks serially); this
module is the multi-GPU form of that loop, not a counterpart of any reference file.
"""
from .encode import EncoderMode, default_context

SHARD_BYTES = 1 << 22   # independent window


def partition(n, world, shard_bytes=SHARD_BYTES):
    """[(offset, length)] per rank: rank g takes shards [g * ceil(K / G), (g + 1) * ceil(K / G)) of the
    K = ceil(n / shard_bytes) shards (SURVEY.md §8e "Partitioning"); trailing ranks may get nothing."""
    if world < 1 or shard_bytes < 64:
        raise ValueError("world >= 1 and shard_bytes >= 64 required")
    k = (n + shard_bytes - 1) // shard_bytes
    per = (k + world - 1) // world if k else 0
    parts = []
    for g in range(world):
        lo = min(n, g * per * shard_bytes)
        hi = min(n, (g + 1) * per * shard_bytes)
        parts.append((lo, hi - lo))
    return parts


def part_roles(parts):
    """(first, last) flag per rank: the first non-empty part opens the stream, the last non-empty one
    closes it; empty parts contribute nothing."""
    live = [g for g, (_, ln) in enumerate(parts) if ln]
    return [(bool(live) and g == live[0], bool(live) and g == live[-1]) for g in range(len(parts))]


def encode_my_part(data, rank, world, quality=11, lgwin=22, mode=EncoderMode.GENERIC, part_encoder=None, shard_bytes=SHARD_BYTES):
    """This rank's compressed part of `data` (bytes-like; only [off - 2, off + len) is read)."""
    mv = memoryview(data).cast("B")
    parts = partition(len(mv), world, shard_bytes)
    off, ln = parts[rank]
    first, last = part_roles(parts)[rank]
    if ln == 0:
        return b""
    prev2 = b"" if first else bytes(mv[off - 2:off])
    enc = part_encoder or _gpu_part_encoder
    return enc(mv[off:off + ln], prev2, first, last, quality, lgwin, mode)


def _gpu_part_encoder(chunk, prev2, first, last, quality, lgwin, mode):
    return default_context().encode_part(chunk, prev2, first, last, quality, lgwin, mode)


def stitch(parts):
    """Host stitcher: parts are byte-aligned, so the stream is their concatenation in rank order."""
    return b"".join(parts)


def brotliEncodeSharded(data, options=None, *, group=None, dst=0, part_encoder=None, shard_bytes=SHARD_BYTES):
    """brotliEncode over every rank of `group` (torch.distributed).  Collective call: every rank passes
    the same input; rank `dst` returns the stream, the others None.  Inputs the reference handles on
    the host (empty, < 64 bytes, quality 0: encode.ts:92-138) are not sharded — rank `dst` encodes them
    alone through brotliEncode."""
    import torch.distributed as dist
    o = dict(options or {})
    q = max(0, min(11, int(o.get("quality", 11))))
    w = max(10, min(24, int(o.get("lgwin", 22))))
    m = o.get("mode", EncoderMode.GENERIC)
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    n = len(memoryview(data).cast("B"))
    if n < 64 or q < 1:
        if rank != dst:
            return None
        from .encode import brotliEncode
        return brotliEncode(data, {"quality": q, "lgwin": w, "mode": m})
    if w < 22:
        shard_bytes = min(shard_bytes, ((1 << w) - 16) & ~3)
    mine = encode_my_part(data, rank, world, q, w, m, part_encoder, shard_bytes)
    gathered = [None] * world if rank == dst else None
    dist.gather_object(mine, gathered, dst=dist.get_global_rank(group, dst) if group is not None else dst, group=group)
    return stitch(gathered) if rank == dst else None
