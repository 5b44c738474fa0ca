"""Multi-GPU MSM by point range: one process per GPU, every rank keeps only its slice
[lo, hi) of a base array resident and receives only its slice of the scalars; the exchange is one all-gather
of the XYZZ partial sums (128 B per rank for G1, 256 B for G2) followed by world-1 point additions on the
host (ugb_g1_sum_xyzz_to_affine).  torch.distributed is plumbing only (NCCL on GPUs, gloo in the CPU tests).
"""
import ctypes

from ._lib import check, lib


def shard_range(n, world, rank):
    """Contiguous slice of [0, n) owned by `rank`; sizes differ by at most one and cover [0, n) exactly."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad world/rank")
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def partial_size(group):
    return 128 if group == 1 else 256


def reduce_partials(partials: bytes, group: int) -> bytes:
    """Sum of XYZZ partial sums -> affine point bytes (host code of the product library, no GPU needed)."""
    sz = partial_size(group)
    if len(partials) % sz:
        raise ValueError("partials length is not a multiple of the XYZZ point size")
    out = ctypes.create_string_buffer(sz // 2)
    fn = lib().ugb_g1_sum_xyzz_to_affine if group == 1 else lib().ugb_g2_sum_xyzz_to_affine
    check(fn(partials, len(partials) // sz, out))
    return out.raw


def allgather_reduce(partial: bytes, group: int, device=None) -> bytes:
    """all-gather this rank's partial sum over the default process group and reduce on every rank."""
    import torch
    import torch.distributed as dist
    sz = partial_size(group)
    if len(partial) != sz:
        raise ValueError("partial sum has the wrong size")
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return reduce_partials(partial, group)
    mine = torch.frombuffer(bytearray(partial), dtype=torch.uint8)
    if device is not None:
        mine = mine.to(device)
    parts = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, mine)
    return reduce_partials(b"".join(bytes(p.cpu().numpy().tobytes()) for p in parts), group)


class ShardedMsm:
    """This rank's slice of one base array, resident on its GPU."""

    def __init__(self, ctx, group, bases_slice):
        self.ctx = ctx
        self.group = group
        self.n = len(bases_slice) // (64 if group == 1 else 128) if not hasattr(bases_slice, "nbytes") else bases_slice.nbytes // (64 if group == 1 else 128)
        self.bases = ctx.upload(bases_slice)

    def partial(self, scalars_dev, n=None, window_bits=0) -> bytes:
        return self.ctx.msm(self.group, self.bases, scalars_dev, self.n if n is None else n, window_bits=window_bits, xyzz=True)
