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


def _allgather_bytes(blob: bytes):
    """All ranks' blobs, in rank order, over the default process group (device tensors under NCCL, host under gloo)."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return [blob]
    mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
    if dist.get_backend() == "nccl":
        mine = mine.cuda()
    parts = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, mine)
    return [bytes(p.cpu().numpy().tobytes()) for p in parts]


class ShardedProver:
    """Point-range sharded Groth16 / UltraGroth prover: this process holds slice `rank` of `world` of every base array
    (ugb_prover_create_sharded, include/ultragroth_b200.h).  `exchange` maps this rank's blob to the list of all ranks'
    blobs; the default is one torch.distributed all-gather.  The proof is assembled on `root` (None elsewhere)."""

    ERR = 512

    def __init__(self, zkey, protocol="groth16", rank=0, world=1, exchange=None, root=0):
        from .prover import _buf, _err, _msg, ProverError
        self.ultra = protocol in ("ultragroth", "ultra_groth")
        self.rank, self.world, self.root = rank, world, root
        self.exchange = exchange or _allgather_bytes
        self.handle = ctypes.c_void_p()
        p, n, keep = _buf(zkey)
        e = _err()
        L = lib()
        L.ugb_prover_partial_size.restype = ctypes.c_uint64
        rc = L.ugb_prover_create_sharded(ctypes.byref(self.handle), p, ctypes.c_ulonglong(n), 1 if self.ultra else 0,
                                         ctypes.c_uint32(rank), ctypes.c_uint32(world), e, ctypes.c_ulonglong(self.ERR - 1))
        if rc:
            self.handle = ctypes.c_void_p()
            raise ProverError(rc, _msg(e))
        self.partial_size = int(L.ugb_prover_partial_size())
        from .prover import proof_size, public_size_for_zkey_buf
        self.proof_cap = proof_size(protocol)
        self.public_cap = public_size_for_zkey_buf(zkey, protocol)

    def close(self):
        if self.handle:
            lib().groth16_prover_destroy(self.handle)
            self.handle = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_blinding(self, r, s, rk=0):
        enc = lambda x: int(x).to_bytes(32, "little")
        lib().ugb_prover_set_blinding(self.handle, enc(r), enc(s), enc(rk))

    # ---- the three phases, usable one by one (tests drive several shards from one process) ----
    def partial(self, wtns, round_partials=None, rk=None) -> bytes:
        from .prover import _buf, _err, _msg, ProverError
        p, n, keep = _buf(wtns)
        e = _err()
        out = ctypes.create_string_buffer(self.partial_size)
        if not self.ultra:
            rc = lib().ugb_groth16_prove_partial(self.handle, p, ctypes.c_ulonglong(n), out, e, ctypes.c_ulonglong(self.ERR - 1))
        else:
            blob = b"".join(round_partials)
            rc = lib().ugb_ultra_final_partial(self.handle, p, ctypes.c_ulonglong(n), blob, ctypes.c_uint32(len(round_partials)),
                                               int(rk).to_bytes(32, "little"), out, e, ctypes.c_ulonglong(self.ERR - 1))
        if rc:
            raise ProverError(rc, _msg(e))
        return out.raw

    def round_partial(self, wtns) -> bytes:
        from .prover import _buf, _err, _msg, ProverError
        p, n, keep = _buf(wtns)
        e = _err()
        out = ctypes.create_string_buffer(128)
        rc = lib().ugb_ultra_round_partial(self.handle, p, ctypes.c_ulonglong(n), out, e, ctypes.c_ulonglong(self.ERR - 1))
        if rc:
            raise ProverError(rc, _msg(e))
        return out.raw

    def assemble(self, partials, wtns=None, rk=None):
        from .prover import _buf, _err, _msg, ProverError
        proof = ctypes.create_string_buffer(self.proof_cap + 1)
        public = ctypes.create_string_buffer(self.public_cap + 1)
        ps, us = ctypes.c_ulonglong(self.proof_cap), ctypes.c_ulonglong(self.public_cap)
        e = _err()
        blob = b"".join(partials)
        if not self.ultra:
            p, n, keep = _buf(wtns)
            rc = lib().ugb_groth16_prove_assemble(self.handle, blob, ctypes.c_uint32(len(partials)), p, ctypes.c_ulonglong(n), proof,
                                                  ctypes.byref(ps), public, ctypes.byref(us), e, ctypes.c_ulonglong(self.ERR - 1))
        else:
            rc = lib().ugb_ultra_prove_assemble(self.handle, blob, ctypes.c_uint32(len(partials)), int(rk).to_bytes(32, "little"), proof,
                                                ctypes.byref(ps), public, ctypes.byref(us), e, ctypes.c_ulonglong(self.ERR - 1))
        if rc:
            raise ProverError(rc, _msg(e))
        return proof.raw.split(b"\0", 1)[0].decode(), public.raw.split(b"\0", 1)[0].decode()

    def prove(self, wtns, rk=None):
        """Collective: every rank calls it with the same witness (and, for UltraGroth, the same r_k < 2^248)."""
        if self.ultra:
            if rk is None:
                raise ValueError("sharded UltraGroth proving needs a common r_k (draw it once and broadcast it)")
            rounds = self.exchange(self.round_partial(wtns))
            parts = self.exchange(self.partial(wtns, rounds, rk))
        else:
            parts = self.exchange(self.partial(wtns))
        if self.root is not None and self.rank != self.root:
            return None
        return self.assemble(parts, wtns=wtns, rk=rk)
