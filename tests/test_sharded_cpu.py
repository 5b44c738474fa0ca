"""CPU test of the N>1 path (world_size 2, gloo): point-range sharding, all-gather of the XYZZ partial sums and
the host reduction of the product library.  The per-rank partial sums come from the oracle here (no GPU); on
the GPU box tests/test_gpu_sharded.py produces them with the CUDA MSM."""
import os
import subprocess
import sys

import pytest

from ultragroth_b200 import sharded

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

def _free_port():
    import socket
    with socket.socket() as so:
        so.bind(("127.0.0.1", 0))
        return so.getsockname()[1]


WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["UGB_ROOT"])
import torch.distributed as dist
from oracle import bn254 as bn, cpu_prover as cp
from tests.util import chain_points, fe_list_bytes, g1_bytes, g1_from_bytes, g2_bytes, g2_from_bytes, scalars_witness_like
from ultragroth_b200 import sharded
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
one = (bn.MONT_R % bn.Q).to_bytes(32, "little")
for group in (1, 2):
    C, gen = (bn.G1, bn.G1_GEN) if group == 1 else (bn.G2, bn.G2_GEN)
    enc, dec = (g1_bytes, g1_from_bytes) if group == 1 else (g2_bytes, g2_from_bytes)
    n = 301
    pts = chain_points(C, gen, n, seed=5)          # same on every rank
    sc = scalars_witness_like(n, seed=6)
    lo, hi = sharded.shard_range(n, world, rank)
    aff = cp.msm(group, b"".join(enc(P) for P in pts[lo:hi]), fe_list_bytes(sc[lo:hi]), hi - lo, threads=1)
    if not any(aff):
        partial = bytes(sharded.partial_size(group))
    else:
        partial = aff + (one * 2 if group == 1 else (one + bytes(32)) * 2)   # ZZ = ZZZ = 1
    got = dec(sharded.allgather_reduce(partial, group))
    want = C.to_affine(C.msm(pts, sc))
    assert got == want, (group, rank)
dist.destroy_process_group()
print("ok", rank)
'''


def test_shard_range_covers_exactly():
    for n in (0, 1, 7, 8, 1000, (1 << 22) + 3):
        for world in (1, 2, 3, 4, 8):
            spans = [sharded.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            chunk = -(-n // world)  # equal chunks (collective-friendly); only the tail may be short or empty
            assert all(hi - lo == chunk or hi == n for lo, hi in spans)
    with pytest.raises(ValueError):
        sharded.shard_range(10, 2, 2)


def test_reduce_partials_identity():
    assert sharded.reduce_partials(bytes(128) * 3, 1) == bytes(64)
    assert sharded.reduce_partials(b"", 2) == bytes(128)
    with pytest.raises(ValueError):
        sharded.reduce_partials(bytes(100), 1)


def test_world2_gloo_sharded_msm(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, UGB_ROOT=ROOT, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    # the rendezvous port is picked free and then released: another process can take it in between, so a failed launch gets ONE more
    # attempt on a fresh port (a wrong result fails both attempts alike)
    for attempt in range(2):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
               "--master-port", str(_free_port()), str(script)]
        res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
        if res.returncode == 0:
            break
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    # the two ranks share one pipe: their "ok <rank>" lines may interleave
    assert res.stdout.count("ok") == 2 and "0" in res.stdout and "1" in res.stdout
