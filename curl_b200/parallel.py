"""Host-side helpers for the multi-GPU path.

The curve layer has no cross-image coupling (the regulariser and the knot gradients are per image),
so it shards along the batch dimension with NO data-path collective: rank r of W processes (one per
GPU, torch.distributed over NCCL/NVLink for the plumbing) owns a contiguous block of images, exactly
what DistributedSampler gives each DDP replica in the reference (main.py:212-213).  The only
collective a training step needs is DDP's gradient all-reduce of the *backbone* weights; the layer
itself has no trainable parameters (colors.py:14-15,22-25,75-76,83-86,184).
"""
import torch
import torch.distributed as dist


def shard_bounds(n_images, rank, world_size):
    """[lo, hi) of the images rank `rank` owns; sizes differ by at most one, earlier ranks larger."""
    if world_size < 1 or not (0 <= rank < world_size) or n_images < 0:
        raise ValueError("bad shard request: n=%d rank=%d world=%d" % (n_images, rank, world_size))
    base, extra = divmod(n_images, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(tensors, rank, world_size):
    """Slices every [B, ...] tensor in `tensors` (None passes through) to this rank's images."""
    n = next(t.shape[0] for t in tensors if t is not None)
    lo, hi = shard_bounds(n, rank, world_size)
    return [None if t is None else t[lo:hi] for t in tensors]


def max_over_ranks(value, device=None):
    """Max of a python float over all ranks (device-timed milliseconds): the slowest rank sets the
    job's time.  No-op without an initialised process group."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_shards(local, world_size, dim=0):
    """all_gather of equally-shaped shards, concatenated along `dim` (used by tests and tools that
    want the whole batch back; the hot path never needs it)."""
    if not (dist.is_available() and dist.is_initialized()) or world_size == 1:
        return local
    parts = [torch.empty_like(local) for _ in range(world_size)]
    dist.all_gather(parts, local.contiguous())
    return torch.cat(parts, dim=dim)
