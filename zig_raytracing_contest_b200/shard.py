"""Host-side sharding of one frame over ranks. One process per GPU, scene replicated.

The reference splits pixels into contiguous ranges over worker threads with no communication
(src/stage3.zig:228-233,247-255); samples of a pixel are independent as well because the RNG is
keyed by absolute (pixel, sample). Two partitions, both without a collective inside the render:

  sample ranges : rank r renders samples [offset, offset+count) of every pixel into a float4
                  accumulator; ONE reduce(SUM) to rank 0, which resolves (1/S, toRGB)
  row bands     : rank r renders rows [y0, y1) and the finished bands are gathered on rank 0
  row strips    : strips of `rows` full rows dealt round-robin (strip i -> rank i % n), so sky rows and
                  geometry rows spread over every rank; the finished strips are gathered on rank 0

Works with any torch.distributed backend (nccl on the GPUs, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def sample_shard(rank, world, total_samples):
    """Contiguous sample range of `rank`: (offset, count). The first total % world ranks get one more."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank %d of %d" % (rank, world))
    base, extra = divmod(int(total_samples), world)
    count = base + (1 if rank < extra else 0)
    offset = rank * base + min(rank, extra)
    return offset, count


def row_band(rank, world, height):
    """Rows [y0, y1) of `rank` — the same split rtc_render applies for tile_rank/tile_world
    (csrc/rtc_api.cu shard_pixels, tile_rows = 0): y0 = h*r/n, y1 = h*(r+1)/n."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank %d of %d" % (rank, world))
    return (height * rank) // world, (height * (rank + 1)) // world


def reduce_accumulators(acc, dst=0, group=None):
    """Sum the per-rank accumulators ([P,4] float32: rgb sum, sample count) onto rank `dst`, in place."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return acc


def gather_row_bands(band, width, height, dst=0, group=None):
    """Assemble the frame from per-rank row bands. `band` is this rank's [y1-y0, width, C] tensor.
    Returns the [height, width, C] frame on rank `dst`, None elsewhere."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        return band
    # bands differ by at most one row: pad to the tallest so one gather of equal tensors does it
    tallest = max(row_band(r, world, height)[1] - row_band(r, world, height)[0] for r in range(world))
    padded = torch.zeros((tallest,) + tuple(band.shape[1:]), dtype=band.dtype, device=band.device)
    padded[:band.shape[0]] = band
    out = [torch.empty_like(padded) for _ in range(world)] if rank == dst else None
    dist.gather(padded, out, dst=dst, group=group)
    if rank != dst:
        return None
    frame = torch.empty((height, width) + tuple(band.shape[2:]), dtype=band.dtype, device=band.device)
    for r in range(world):
        y0, y1 = row_band(r, world, height)
        frame[y0:y1] = out[r][:y1 - y0]
    return frame


def strips_of(rank, world, height, rows):
    """Row ranges [(y0, y1), ...] of `rank` when strips of `rows` rows are dealt round-robin — the split
    rtc_render applies for tile_rows > 0 (csrc/rtc_api.cu shard_pixels)."""
    if world < 1 or not 0 <= rank < world or rows < 1:
        raise ValueError("rank %d of %d, %d rows per strip" % (rank, world, rows))
    nstrips = (height + rows - 1) // rows
    return [(k * rows, min(height, (k + 1) * rows)) for k in range(rank, nstrips, world)]


def gather_strips(frame, rows, dst=0, group=None):
    """Assemble the frame from per-rank strips with ONE gather. `frame` is [H, W, C] on every rank with
    this rank's strips filled in; returns the complete frame on rank `dst`, None elsewhere."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        return frame
    frame = frame.contiguous()                    # the strip views below must alias ONE buffer (reshape of a strided frame copies)
    H = frame.shape[0]
    nstrips = (H + rows - 1) // rows
    per_rank = (nstrips + world - 1) // world
    row_elems = int(frame[0].numel())
    padded = frame
    if nstrips * rows != H:                       # the last strip is short: pad so strips are equal-sized
        padded = torch.zeros((nstrips * rows,) + tuple(frame.shape[1:]), dtype=frame.dtype, device=frame.device)
        padded[:H] = frame
    view = padded.reshape(nstrips, rows * row_elems)
    mine = view[rank::world]
    compact = torch.zeros((per_rank, rows * row_elems), dtype=frame.dtype, device=frame.device)
    compact[:mine.shape[0]] = mine
    out = [torch.empty_like(compact) for _ in range(world)] if rank == dst else None
    dist.gather(compact, out, dst=dst, group=group)
    if rank != dst:
        return None
    for r in range(world):
        n_r = len(range(r, nstrips, world))
        view[r::world] = out[r][:n_r]
    return padded[:H]


def broadcast_comm_id(make_id, src=0, group=None):
    """The library's NCCL communicator (rtc_comm_create) needs the same 128-byte id on every rank: rank `src` makes it
    (make_id = rtc.comm_unique_id) and torch.distributed — whatever its backend — hands the bytes around. This is all
    the host framework contributes to the data path of a sharded frame; the collective itself is issued by librtc.so."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return make_id()
    box = [make_id() if dist.get_rank(group) == src else None]
    dist.broadcast_object_list(box, src=src, group=group)
    if not isinstance(box[0], (bytes, bytearray)) or len(box[0]) != 128:
        raise RuntimeError("communicator id did not arrive (%r)" % type(box[0]))
    return bytes(box[0])


def max_over_ranks(value, device=None, group=None):
    """A timing as the contract wants it: the MAX over the ranks (float in, float out)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t[0])
