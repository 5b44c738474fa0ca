"""GPU: synthetic circuits with a valid trusted setup (harness tooling, ugb_synth_build) and the proofs made
from them -- Groth16 and UltraGroth through the reference's C API, bit-exact against the oracle at small
sizes, accepted by the repo's verifier equation at every size."""
import json

import pytest

from oracle import bn254 as bn
from oracle import groth16 as og
from tests.util import g1_from_bytes, g2_from_bytes
from ultragroth_b200 import prover as up
from ultragroth_b200 import verifier as uv

pytestmark = pytest.mark.gpu


def test_fixed_base_mul_vs_oracle(ctx):
    ks = [0, 1, 2, 255, 256, 65537, bn.R - 1, bn.R - 2, (1 << 200) + 12345, 0x1234567890abcdef << 120]
    out1 = ctx.fixed_base_mul(1, ks)
    out2 = ctx.fixed_base_mul(2, ks)
    for i, k in enumerate(ks):
        assert g1_from_bytes(out1[64 * i:64 * i + 64]) == (bn.G1.mul_affine(bn.G1_GEN, k) if k else None)
        assert g2_from_bytes(out2[128 * i:128 * i + 128]) == (bn.G2.mul_affine(bn.G2_GEN, k) if k else None)


@pytest.mark.parametrize("ultra,logn,lookup_bits,chunks", [(False, 4, 0, 0), (False, 12, 0, 0), (True, 7, 3, 20), (True, 13, 7, 2048)])
def test_gpu_and_cpu_generators_emit_the_same_files(ctx, ultra, logn, lookup_bits, chunks):
    """The GPU generator (ugb_synth_build, the fixed-base multiplications on the device) and the oracle's CPU generator
    (oracle/cpu_synth.cpp, used by bench.py's CPU arm so that it never loads the product library) must produce identical bytes:
    zkey, witness and verification key."""
    from oracle import cpu_prover as cp
    gz, gw, gv = ctx.synth(logn, seed=4242 + logn, ultra=ultra, lookup_bits=lookup_bits, n_chunks=chunks)
    cz, cw, cv = cp.synth(logn, seed=4242 + logn, ultra=ultra, lookup_bits=lookup_bits, n_chunks=chunks)
    assert gz.tobytes() == cz.tobytes()
    assert gw.tobytes() == cw.tobytes()
    assert gv == cv


@pytest.mark.parametrize("logn", [2, 5, 8])
def test_synth_groth16_small_bit_exact(ctx, logn):
    zk, wt, vkj = ctx.synth(logn, seed=100 + logn)
    zk, wt = zk.tobytes(), wt.tobytes()
    vk = og.vk_from_snarkjs_json(vkj)
    r, s = 987654321987654321, (1 << 240) + 77
    want_proof, want_public, inter = og.groth16_prove(zk, wt, r, s, want_intermediates=True)
    # the generated key is a valid setup: the ORACLE's proof verifies against the generated vk
    assert og.groth16_verify(vk, want_proof, want_public)
    p = up.Prover(zk)
    p.set_blinding(r, s)
    pj, uj = p.prove(wt)
    h = p.last_h(1 << logn)
    assert [int.from_bytes(h[i:i + 32], "little") for i in range(0, len(h), 32)] == inter["h"]
    assert pj == og.proof_to_json(want_proof)
    assert uj == og.public_to_json(want_public)
    p.close()


@pytest.mark.parametrize("logn,lookup_bits,chunks", [(6, 2, 5), (9, 4, 100)])
def test_synth_ultragroth_small_bit_exact(ctx, logn, lookup_bits, chunks):
    zk, wt, vkj = ctx.synth(logn, seed=300 + logn, ultra=True, lookup_bits=lookup_bits, n_chunks=chunks)
    zk, wt = zk.tobytes(), wt.tobytes()
    vk = og.vk_from_ultragroth_json(vkj)
    rk, r, s = (1 << 247) - 3, 1234567, (1 << 200) + 9
    want_proof, want_public, inter = og.ultragroth_prove(zk, wt, rk, r, s, want_intermediates=True)
    assert og.ultragroth_verify(vk, want_proof, want_public)
    p = up.Prover(zk, "ultragroth")
    p.set_blinding(r, s, rk)
    pj, uj = p.prove(wt)
    h = p.last_h(1 << logn)
    assert [int.from_bytes(h[i:i + 32], "little") for i in range(0, len(h), 32)] == inter["h"]
    assert pj == og.proof_to_json(want_proof)
    assert uj == og.public_to_json(want_public)
    # random blinding: still verifies; tampered public input fails
    p.set_blinding(None, None)
    pj2, uj2 = p.prove(wt)
    public = [int(x) for x in json.loads(uj2)]
    assert og.ultragroth_verify(vk, og.proof_from_json(pj2), public)
    assert not og.ultragroth_verify(vk, og.proof_from_json(pj2), [(public[0] + 1) % bn.R])
    p.close()


@pytest.mark.parametrize("logn", [14, 18])
def test_synth_groth16_large_verifies(ctx, logn):
    zk, wt, vkj = ctx.synth(logn, seed=7)
    vk = og.vk_from_snarkjs_json(vkj)
    p = up.Prover(zk)
    pj, uj = p.prove(wt)
    public = [int(x) for x in json.loads(uj)]
    assert og.groth16_verify(vk, og.proof_from_json(pj), public)
    assert not og.groth16_verify(vk, og.proof_from_json(pj), [(public[0] - 1) % bn.R])
    # and under the product's own verifier C API (include/verifier.h)
    assert uv.groth16_verify(pj, uj, vkj) == (uv.VERIFIER_VALID_PROOF, "")
    assert uv.groth16_verify(pj, json.dumps([str((public[0] - 1) % bn.R)]), vkj)[0] == uv.VERIFIER_INVALID_PROOF
    p.close()


def test_synth_ultragroth_large_verifies(ctx):
    zk, wt, vkj = ctx.synth(17, seed=8, ultra=True, lookup_bits=10, n_chunks=30000)
    vk = og.vk_from_ultragroth_json(vkj)
    p = up.Prover(zk, "ultragroth")
    pj, uj = p.prove(wt)
    public = [int(x) for x in json.loads(uj)]
    assert og.ultragroth_verify(vk, og.proof_from_json(pj), public)
    assert uv.ultra_groth_verify(pj, uj, vkj) == (uv.VERIFIER_VALID_PROOF, "")
    p.close()


def test_config1_groth16_2pow20_full_size(ctx):
    """BASELINE.json configs[1]: synthetic Groth16 circuit, 2^20 constraints, one B200.  At this size the oracle prover is
    out of reach, so parity is through the verifier equation (product C++ verifier and oracle verifier), a tampered input
    must fail, and the resident prover must give a fresh valid proof on every call."""
    zk, wt, vkj = ctx.synth(20, seed=0xB200 + 20)
    p = up.Prover(zk)
    del zk
    pj, uj = p.prove(wt)
    public = [int(x) for x in json.loads(uj)]
    assert uv.groth16_verify(pj, uj, vkj) == (uv.VERIFIER_VALID_PROOF, "")
    assert og.groth16_verify(og.vk_from_snarkjs_json(vkj), og.proof_from_json(pj), public)
    assert uv.groth16_verify(pj, json.dumps([str((public[0] + 1) % bn.R)]), vkj)[0] == uv.VERIFIER_INVALID_PROOF
    pj2, uj2 = p.prove(wt)
    assert pj2 != pj and uj2 == uj
    assert uv.groth16_verify(pj2, uj2, vkj) == (uv.VERIFIER_VALID_PROOF, "")
    p.close()
