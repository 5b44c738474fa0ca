"""GPU: sharded proving.  Every shard holds 1/world of each base array (H and the round bases by contiguous range, the
witness-paired bases interleaved: signal s on shard s % world -- world 3 and 5 make the shares ragged) and returns
partial MSM sums; the assembled proof must be BYTE-IDENTICAL to the unsharded prover's for the same blinding, at every
world size (the shard sums are a re-association of the same group sum, and the proof is an affine, canonical value).

Shards are driven from one process here (world 2, 3 and 5 on one GPU); test_two_process_gloo runs the real collective
path (torch.distributed all-gather, gloo, both ranks on cuda:0 because NCCL refuses two ranks per device)."""
import json
import os
import subprocess
import sys

import pytest

from oracle import groth16 as og
from ultragroth_b200 import prover as up
from ultragroth_b200 import sharded

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R_FIX, S_FIX, RK_FIX = 0x1234567890abcdef1234567890abcdef1234567890abcdef1234567890ab, (1 << 247) + 12345, (1 << 200) + 99


def _prove_sharded(zk, wt, protocol, world, rk=None):
    shards = [sharded.ShardedProver(zk, protocol, rank=r, world=world, exchange=lambda b: [b], root=None) for r in range(world)]
    for s in shards:
        s.set_blinding(R_FIX, S_FIX, RK_FIX)
    if protocol == "groth16":
        parts = [s.partial(wt) for s in shards]
        out = [s.assemble(parts, wtns=wt) for s in (shards[0], shards[-1])]
    else:
        rounds = [s.round_partial(wt) for s in shards]
        parts = [s.partial(wt, rounds, rk) for s in shards]
        out = [s.assemble(parts, rk=rk) for s in (shards[0], shards[-1])]
    for s in shards:
        s.close()
    assert out[0] == out[1]  # any rank can assemble
    return out[0]


@pytest.mark.parametrize("world", [2, 3, 5])
def test_groth16_sharded_equals_unsharded(ctx, world):
    zk, wt, vkj = ctx.synth(10, seed=4242)
    zk, wt = zk.tobytes(), wt.tobytes()
    p = up.Prover(zk)
    p.set_blinding(R_FIX, S_FIX)
    want = p.prove(wt)
    p.close()
    got = _prove_sharded(zk, wt, "groth16", world)
    assert got == want
    vk = og.vk_from_snarkjs_json(vkj)
    assert og.groth16_verify(vk, og.proof_from_json(got[0]), [int(x) for x in json.loads(got[1])])


def test_groth16_sharded_reference_fixture():
    gold = os.path.join(os.path.dirname(__file__), "golden", "ref_testdata")
    zk = open(os.path.join(gold, "circuit_final.zkey"), "rb").read()
    wt = open(os.path.join(gold, "witness.wtns"), "rb").read()
    want_proof, want_public = og.groth16_prove(zk, wt, R_FIX, S_FIX)
    got = _prove_sharded(zk, wt, "groth16", 4)
    assert got[0] == og.proof_to_json(want_proof) and got[1] == og.public_to_json(want_public)


@pytest.mark.parametrize("world", [2, 3])
def test_ultragroth_sharded_equals_unsharded(ctx, world):
    zk, wt, vkj = ctx.synth(10, seed=777, ultra=True, lookup_bits=4, n_chunks=256)
    zk, wt = zk.tobytes(), wt.tobytes()
    p = up.Prover(zk, "ultragroth")
    p.set_blinding(R_FIX, S_FIX, RK_FIX)
    want = p.prove(wt)
    p.close()
    got = _prove_sharded(zk, wt, "ultragroth", world, rk=RK_FIX)
    assert got == want
    proof = og.proof_from_json(got[0])
    public = [int(x) for x in json.loads(got[1])] if got[1] != "null" else []
    assert og.ultragroth_verify(og.vk_from_ultragroth_json(vkj), proof, public)


@pytest.mark.parametrize("protocol,world", [("groth16", 2), ("groth16", 3), ("groth16", 5), ("ultragroth", 2), ("ultragroth", 4)])
def test_phasewise_protocol_equals_unsharded(ctx, protocol, world):
    """The de-replicated protocol (ugb_shard_*): witness SLICES uploaded and all-gathered, the a / b / c transforms on different
    shards, every shard forming only its range of h -- byte-identical proofs to the unsharded prover (shards in one process, the
    exchanges as device-to-device copies: sharded.prove_local)."""
    ultra = protocol == "ultragroth"
    if ultra:
        zk, wt, vkj = ctx.synth(10, seed=779, ultra=True, lookup_bits=4, n_chunks=256)
    else:
        zk, wt, vkj = ctx.synth(10, seed=4243)
    zk, wt = zk.tobytes(), wt.tobytes()
    p = up.Prover(zk, protocol)
    p.set_blinding(R_FIX, S_FIX, RK_FIX)
    want = p.prove(wt)
    p.close()
    shards = [sharded.ShardedProver(zk, protocol, rank=r, world=world, exchange=lambda b: [b], root=None) for r in range(world)]
    for s in shards:
        s.set_blinding(R_FIX, S_FIX, RK_FIX)
    out = sharded.prove_local(shards, wt, rk=RK_FIX if ultra else None)
    for s in shards:
        s.close()
    assert out[0] == out[1] == want


def _free_port():
    import socket
    with socket.socket() as so:
        so.bind(("127.0.0.1", 0))
        return so.getsockname()[1]


WORKER = r'''
import os, sys, json
sys.path.insert(0, os.environ["UGB_ROOT"])
import torch.distributed as dist
import ultragroth_b200 as ug
from ultragroth_b200 import sharded, prover as up
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
ctx = ug.Context(0)
zk, wt, vkj = ctx.synth(9, seed=31337)          # deterministic: the same key on every rank
zk, wt = zk.tobytes(), wt.tobytes()
sp = sharded.ShardedProver(zk, "groth16", rank=rank, world=world)
sp.set_blinding(11, 22)
res = sp.prove(wt)
if rank == 0:
    p = up.Prover(zk); p.set_blinding(11, 22)
    assert res == p.prove(wt), "sharded proof differs from the unsharded one"
    p.close()
else:
    assert res is None
sp.close()
dist.destroy_process_group()
print("ok", rank)
'''


def test_two_process_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, UGB_ROOT=ROOT, MASTER_ADDR="127.0.0.1")
    for attempt in range(2):  # the free port is released before torchrun binds it: one more attempt on a fresh port if the launch fails
        out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                              "--master-port", str(_free_port()), str(script)], env=env, capture_output=True, text=True, timeout=600)
        if out.returncode == 0:
            break
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    # the two ranks share one pipe: their "ok <rank>" lines may interleave character by character
    assert out.stdout.count("ok") == 2 and "0" in out.stdout and "1" in out.stdout
