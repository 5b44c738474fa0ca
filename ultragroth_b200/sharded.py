"""Multi-GPU MSM by point range: one process per GPU, every rank keeps only its slice
[lo, hi) of a base array resident and receives only its slice of the scalars; the exchange is one all-gather
of the XYZZ partial sums (128 B per rank for G1, 256 B for G2) followed by world-1 point additions on the
host (ugb_g1_sum_xyzz_to_affine).  torch.distributed is plumbing only (NCCL on GPUs, gloo in the CPU tests).
That is the standalone MSM (ShardedMsm, sweep.py).  Inside a ShardedProver only H, the round bases and the witness UPLOAD are cut
into such contiguous ranges; the witness-paired bases (A, B1, B2, C) are dealt out interleaved, signal s on shard s % world, so that
every shard sees the same mix of small and full-width scalars.
"""
import ctypes

from ._lib import check, lib


def shard_range(n, world, rank):
    """Contiguous slice of [0, n) owned by `rank`: equal chunks of ceil(n / world) (the last ones may be short or empty), covering
    [0, n) exactly -- equal chunks so that slices can travel in equal-sized all-gather / scatter collectives."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad world/rank")
    chunk = -(-n // world)
    lo = min(n, rank * chunk)
    return lo, min(n, lo + chunk)


class ShardLayout(ctypes.Structure):
    """ugb_shard_layout (include/ultragroth_b200.h)"""
    _fields_ = [("w", ctypes.c_void_p), ("a", ctypes.c_void_p), ("b", ctypes.c_void_p), ("c", ctypes.c_void_p),
                ("nVars", ctypes.c_uint64), ("domainSize", ctypes.c_uint64), ("w_chunk", ctypes.c_uint64), ("h_chunk", ctypes.c_uint64),
                ("rank", ctypes.c_uint32), ("world", ctypes.c_uint32), ("device", ctypes.c_int), ("stream", ctypes.c_void_p)]


def vector_owner(v, world):
    """Which rank transforms vector v (0: a, 1: b, 2: c) of the H pipeline (src/groth16.cpp:110-140: three independent transforms).
    world >= 3: ranks 0, 1, 2; world == 2: rank 0 takes a and b, rank 1 takes c; world == 1: rank 0 takes all."""
    if world >= 3:
        return v
    if world == 2:
        return 1 if v == 2 else 0
    return 0


def vector_mask(rank, world):
    return sum(1 << v for v in range(3) if vector_owner(v, world) == rank)


class _RawCuda:
    """CUDA array interface over a raw device pointer (so torch can alias the prover's own buffers, no copies)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def device_view(ptr, nbytes, device):
    import torch
    return torch.as_tensor(_RawCuda(ptr, nbytes), device=torch.device("cuda", device))


class DistComm:
    """The exchanges of ShardedProver.prove2 over torch.distributed (NCCL, NVLink / NVSwitch) on tensors that ALIAS the prover's device
    buffers, enqueued against the prover's own stream -- no staging copies and no host synchronisation:
      * all_gather_witness: in-place all-gather of the equal-sized witness slices (every rank ends with the whole witness in HBM
        although only 1 / world of it crossed PCIe on each rank);
      * scatter_vectors: the owner of a', b', c' sends rank j its point range; one batched group of NCCL send / recv, asynchronous,
        so the witness MSMs run while the slices travel."""

    def __init__(self, sp):
        import torch
        self.torch = torch
        self.sp = sp
        lay = sp.layout()
        self.lay = lay
        dev = lay.device
        self.stream = torch.cuda.ExternalStream(lay.stream, device=torch.device("cuda", dev))
        self.w = device_view(lay.w, lay.w_chunk * lay.world * 32, dev)
        self.vec = [device_view(p, lay.h_chunk * lay.world * 32, dev) for p in (lay.a, lay.b, lay.c)]

    def stream_scope(self):
        return self.torch.cuda.stream(self.stream)

    def all_gather_witness(self):
        import torch.distributed as dist
        lay = self.lay
        if lay.world == 1:
            return
        cb = lay.w_chunk * 32
        dist.all_gather_into_tensor(self.w, self.w[lay.rank * cb:(lay.rank + 1) * cb])

    def scatter_vectors(self):
        import torch.distributed as dist
        lay = self.lay
        if lay.world == 1:
            return []
        cb = lay.h_chunk * 32
        ops = []
        for v in range(3):
            owner = vector_owner(v, lay.world)
            if lay.rank == owner:
                for j in range(lay.world):
                    if j != owner:
                        ops.append(dist.P2POp(dist.isend, self.vec[v][j * cb:(j + 1) * cb], j))
            else:
                ops.append(dist.P2POp(dist.irecv, self.vec[v][lay.rank * cb:(lay.rank + 1) * cb], owner))
        return dist.batch_isend_irecv(ops) if ops else []

    def wait(self, pending):
        for req in pending:
            req.wait()


def prove_local(shards, wtns, rk=None):
    """The prove2 protocol over shards that live in THIS process (one GPU holding several shards in the tests, or one process driving
    several GPUs): the exchanges are plain device-to-device copies between the shards' buffers."""
    import torch
    world = len(shards)
    lays = [s.layout() for s in shards]
    W = [device_view(l.w, l.w_chunk * l.world * 32, l.device) for l in lays]
    V = [[device_view(p, l.h_chunk * l.world * 32, l.device) for p in (l.a, l.b, l.c)] for l in lays]
    streams = [torch.cuda.ExternalStream(l.stream, device=torch.device("cuda", l.device)) for l in lays]

    def sync_all():
        for st in streams:
            st.synchronize()
        torch.cuda.synchronize()

    for s in shards:
        s.upload_slice(wtns)
    sync_all()
    cb = lays[0].w_chunk * 32
    for src in range(world):
        for dst in range(world):
            if dst != src:
                W[dst][src * cb:(src + 1) * cb].copy_(W[src][src * cb:(src + 1) * cb])
    sync_all()
    ultra = shards[0].ultra
    if ultra:
        rounds = [s.round_partial_resident() for s in shards]
        for s in shards:
            s.lookup(rounds, rk)
    for r, s in enumerate(shards):
        s.h_transform(vector_mask(r, world))
    sync_all()
    hb = lays[0].h_chunk * 32
    for v in range(3):
        owner = vector_owner(v, world)
        for dst in range(world):
            if dst != owner:
                V[dst][v][dst * hb:(dst + 1) * hb].copy_(V[owner][v][dst * hb:(dst + 1) * hb])
    sync_all()
    for s in shards:
        s.witness_msms()
    for s in shards:
        s.h_msm()
    parts = [s.finish() for s in shards]
    return [s.assemble(parts, wtns=wtns, rk=rk) for s in (shards[0], shards[-1])]


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
    world = dist.get_world_size()
    if dist.get_backend() == "nccl":
        # one flat output and ONE device -> host read (a list of `world` tensors costs `world` small copies per exchange)
        mine = mine.cuda()
        flat = torch.empty(world * mine.numel(), dtype=torch.uint8, device=mine.device)
        dist.all_gather_into_tensor(flat, mine)
        raw = flat.cpu().numpy().tobytes()
        n = len(blob)
        return [raw[i * n:(i + 1) * n] for i in range(world)]
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    return [bytes(p.numpy().tobytes()) for p in parts]


class ShardedProver:
    """Point-sharded Groth16 / UltraGroth prover: this process holds share `rank` of `world` of every base array
    (ugb_prover_create_sharded, include/ultragroth_b200.h: H and the round bases by contiguous range, the witness-paired bases
    interleaved, signal s on shard s % world).  `exchange` maps this rank's blob to the list of all ranks'
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

    # ---- phase-by-phase protocol (ugb_shard_*, include/ultragroth_b200.h): nothing replicated that does not have to be ----
    def layout(self):
        lay = ShardLayout()
        check(lib().ugb_shard_layout_get(self.handle, ctypes.byref(lay)))
        return lay

    def _phase(self, name, *args):
        from .prover import _err, _msg, ProverError
        e = _err()
        rc = getattr(lib(), name)(self.handle, *args, e, ctypes.c_ulonglong(self.ERR - 1))
        if rc:
            raise ProverError(rc, _msg(e))

    def upload_slice(self, wtns):
        from .prover import _buf
        p, n, keep = _buf(wtns)
        self._phase("ugb_shard_upload_slice", p, ctypes.c_ulonglong(n))

    def round_partial_resident(self) -> bytes:
        out = ctypes.create_string_buffer(128)
        self._phase("ugb_shard_round_partial", out)
        return out.raw

    def lookup(self, round_partials, rk):
        self._phase("ugb_shard_lookup", b"".join(round_partials), ctypes.c_uint32(len(round_partials)), int(rk).to_bytes(32, "little"))

    def h_transform(self, mask):
        self._phase("ugb_shard_h_transform", ctypes.c_int(mask))

    def witness_msms(self):
        self._phase("ugb_shard_witness_msms")

    def h_msm(self):
        self._phase("ugb_shard_h_msm")

    def finish(self) -> bytes:
        out = ctypes.create_string_buffer(self.partial_size)
        self._phase("ugb_shard_finish", out)
        return out.raw

    def prove2(self, wtns, rk=None, comm=None, trace=None):
        """Collective, phase by phase: this rank uploads only its witness slice, the slices are all-gathered between the GPUs, the three
        H-pipeline transforms run on up to three ranks and each rank receives its point range of a', b', c' (see DistComm).  Every rank
        passes the same witness file (only rank `root` reads the public signals from it) and, for UltraGroth, the same r_k.
        trace: a dict that receives this rank's wall-clock milliseconds per phase; the stream is drained at every seam to attribute them,
        so a traced proof is slower than an untraced one and its phases do not overlap (a diagnosis aid, never a bench number)."""
        import time
        comm = comm or DistComm(self)
        t_last = [time.perf_counter()]

        def seam(name):
            if trace is None:
                return
            comm.stream.synchronize()
            now = time.perf_counter()
            trace[name] = trace.get(name, 0.0) + (now - t_last[0]) * 1e3
            t_last[0] = now

        with comm.stream_scope():
            self.upload_slice(wtns)
            seam("upload_slice")
            comm.all_gather_witness()
            seam("all_gather_witness")
            if self.ultra:
                if rk is None:
                    raise ValueError("sharded UltraGroth proving needs a common r_k (draw it once and broadcast it)")
                rounds = self.exchange(self.round_partial_resident())
                seam("round_msm_and_exchange")
                self.lookup(rounds, rk)
                seam("lookup")
            self.h_transform(vector_mask(self.rank, self.world))
            seam("h_transform")
            pending = comm.scatter_vectors()
            self.witness_msms()
            seam("witness_msms")
            comm.wait(pending)
            seam("scatter_wait")
            self.h_msm()
            mine = self.finish()
            seam("h_msm_and_finish")
        parts = self.exchange(mine)
        seam("exchange_partials")
        if self.root is not None and self.rank != self.root:
            return None
        out = self.assemble(parts, wtns=wtns, rk=rk)
        seam("assemble")
        return out

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
