"""Multi-GPU plumbing for the paths that shard: the batched channelizer shards BY CHANNEL — channels are
independent receivers, no state is shared — so there is no data-path collective; the only exchange is ONE gather of
the demodulated audio (or spectra) onto rank 0 per batch. One process per GPU, torch.distributed (NCCL over
NVLink/NVSwitch on the GPU box, gloo in the CPU tests). Nothing here touches sample arithmetic."""
import torch
import torch.distributed as dist


def shard_range(units, world, rank):
    """Contiguous block partition of `units` (channels / streams) over `world` ranks: the first units % world ranks
    get one extra. Returns (first, end)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    base, extra = divmod(units, world)
    first = rank * base + min(rank, extra)
    return first, first + base + (1 if rank < extra else 0)


def channel_frequencies(nch, first_centre=-30.69e6, spacing=60e3):
    """BASELINE config 4 centre frequencies: -30.69 MHz + k * 60 kHz."""
    return [first_centre + spacing * k for k in range(nch)]


def gather_rows(local, units, dst=0):
    """Gather row-sharded results onto `dst`. local: [rows_of_this_rank, m] tensor (complex128 or real) laid out as
    shard_range() assigns rows. Returns the [units, m] tensor on dst, None elsewhere. Single process: returns local.
    Uneven shards are padded to the largest shard for the collective and trimmed afterwards."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    world, rank = dist.get_world_size(), dist.get_rank()
    is_cpx = local.is_complex()
    buf = torch.view_as_real(local) if is_cpx else local
    rows_max = shard_range(units, world, 0)[1]
    pad = rows_max - buf.shape[0]
    if pad:
        buf = torch.cat([buf, buf.new_zeros((pad,) + tuple(buf.shape[1:]))], dim=0)
    buf = buf.contiguous()
    if dist.get_backend() == "nccl":
        # one all_gather_into_tensor over NVLink (a gather to one root has no faster NCCL primitive); rank dst keeps it
        full = buf.new_empty((world * rows_max,) + tuple(buf.shape[1:]))
        dist.all_gather_into_tensor(full, buf)
        parts = list(full.split(rows_max, dim=0)) if rank == dst else None
    else:
        parts = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
        dist.gather(buf, parts, dst=dst)
    if rank != dst:
        return None
    out = torch.cat([p[: shard_range(units, world, r)[1] - shard_range(units, world, r)[0]] for r, p in enumerate(parts)], dim=0)
    return torch.view_as_complex(out) if is_cpx else out


def max_over_ranks(value, device):
    """Device-timed milliseconds -> max over ranks (the contract's timing rule)."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])
