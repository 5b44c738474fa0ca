"""GPU parity: Groth16 through the reference's C API (include/prover.h) vs the oracle, on the reference's
own fixture (testdata/circuit_final.zkey + witness.wtns, committed under tests/golden/ref_testdata)."""
import json
import os

import pytest

from oracle import bn254 as bn
from oracle import groth16 as og
from ultragroth_b200 import prover as up

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_testdata")


@pytest.fixture(scope="module")
def fixture():
    zk = open(os.path.join(GOLD, "circuit_final.zkey"), "rb").read()
    wt = open(os.path.join(GOLD, "witness.wtns"), "rb").read()
    vk = og.vk_from_snarkjs_json(open(os.path.join(GOLD, "verification_key.json")).read())
    return zk, wt, vk


def test_groth16_testdata_bit_exact(fixture):
    zk, wt, vk = fixture
    r, s = 0x1234567890abcdef1234567890abcdef1234567890abcdef1234567890ab, (1 << 247) + 12345
    want_proof, want_public, inter = og.groth16_prove(zk, wt, r, s, want_intermediates=True)
    p = up.Prover(zk, "groth16")
    p.set_blinding(r, s)
    proof_json, public_json = p.prove(wt)
    # H coefficients bit-exact
    h = p.last_h(1024)
    got_h = [int.from_bytes(h[i:i + 32], "little") for i in range(0, len(h), 32)]
    assert got_h == inter["h"]
    assert proof_json == og.proof_to_json(want_proof)
    assert public_json == og.public_to_json(want_public)
    # and it verifies under the repo's verifier equation; a tampered public input must fail (CI test of the reference)
    proof = og.proof_from_json(proof_json)
    public = [int(x) for x in json.loads(public_json)]
    assert og.groth16_verify(vk, proof, public)
    assert not og.groth16_verify(vk, proof, [(public[0] - 1) % bn.R])
    # second proof on the same resident prover, random blinding: differs but verifies
    p.set_blinding(None, None)
    pj2, _ = p.prove(wt)
    assert pj2 != proof_json
    assert og.groth16_verify(vk, og.proof_from_json(pj2), public)
    p.close()


def test_groth16_oneshot_and_errors(fixture):
    zk, wt, vk = fixture
    pj, uj = up.prove_oneshot(zk, wt, "groth16")
    assert og.groth16_verify(vk, og.proof_from_json(pj), [int(x) for x in json.loads(uj)])
    assert up.proof_size("groth16") == 810 and up.proof_size("ultragroth") == 1400
    assert up.public_size_for_zkey_buf(zk) == 1 * 82 + 4
    with pytest.raises(up.ProverError) as e:
        up.Prover(b"zkeX" + zk[4:])
    assert e.value.code == up.PROVER_ERROR and "Invalid file type" in e.value.msg
    with pytest.raises(up.ProverError) as e:
        up.Prover(zk, "ultragroth")
    assert "zkey file is not ultragroth" in e.value.msg
    p = up.Prover(zk)
    with pytest.raises(up.ProverError) as e:
        p.prove(wt, proof_capacity=100)
    assert e.value.code == up.PROVER_ERROR_SHORT_BUFFER and "Proof buffer is too short. Minimum size: 810, actual size: 100" in e.value.msg
    with pytest.raises(up.ProverError) as e:
        p.prove(wt, public_capacity=10)
    assert e.value.code == up.PROVER_ERROR_SHORT_BUFFER
    # witness with the wrong number of signals
    from oracle.formats import Wtns, write_wtns
    sig = Wtns(wt).signals()
    with pytest.raises(up.ProverError) as e:
        p.prove(write_wtns(sig[:-1]))
    assert e.value.code == up.PROVER_INVALID_WITNESS_LENGTH
    assert e.value.msg == "Invalid witness length. Circuit: 1003, witness: 1002"
    with pytest.raises(up.ProverError) as e:
        p.prove(b"wtnX" + wt[4:])
    assert e.value.code == up.PROVER_ERROR
    p.close()


def test_groth16_matches_committed_golden_proof(fixture):
    """The CUDA prover against the committed golden proof (tests/golden/testdata_proof.json): no oracle run involved."""
    import hashlib
    zk, wt, vk = fixture
    g = json.load(open(os.path.join(os.path.dirname(GOLD), "testdata_proof.json")))
    p = up.Prover(zk, "groth16")
    p.set_blinding(int(g["r"], 16), int(g["s"], 16))
    pj, uj = p.prove(wt)
    assert json.loads(pj) == g["proof"] and json.loads(uj) == g["public"]
    assert hashlib.sha256(p.last_h(g["h_count"])).hexdigest() == g["h_sha256"]
    p.close()
    # the one-shot entry point (no window tables) gives the same verdict
    pj1, uj1 = up.prove_oneshot(zk, wt, "groth16")
    assert og.groth16_verify(vk, og.proof_from_json(pj1), [int(x) for x in json.loads(uj1)])
