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


@pytest.mark.parametrize("protocol", ["groth16", "ultragroth"])
def test_pipelined_batch_equals_sequential_proofs(ctx, protocol):
    """ugb_prover_prove_batch (N3): witness k+1 uploads on a copy stream into the second witness buffer while proof k computes and the
    host tail of proof k is deferred behind proof k+1 -- with fixed blinding every proof must equal the one *_prover_prove gives
    for the same witness.  The witnesses differ from each other (a few perturbed signals: the prover does not check satisfiability,
    the proof points are still a deterministic function of the witness), so a buffer mix-up cannot go unnoticed."""
    ultra = protocol == "ultragroth"
    if ultra:
        zk, wt, _ = ctx.synth(11, seed=555, ultra=True, lookup_bits=5, n_chunks=400)
    else:
        zk, wt, _ = ctx.synth(11, seed=556)
    zk, wt = zk.tobytes(), wt.tobytes()
    from oracle import formats as fm
    off2 = fm.BinFile(wt, "wtns", 2).sections[2][0][0]
    variants = []
    for k in range(5):
        b = bytearray(wt)
        for sig in (7 + k, 100 + 3 * k, 900):
            b[off2 + 32 * sig:off2 + 32 * sig + 8] = (0x1234567 * (k + 1) + sig).to_bytes(8, "little")
        variants.append(bytes(b))
    p = up.Prover(zk, protocol)
    p.set_blinding(1111, 2222, 3333)
    want = [p.prove(w) for w in variants]
    assert len({w[0] for w in want}) == len(variants)
    got = up.prove_batch(p, variants)
    assert got == want
    # and again on the same object (buffers are left in a consistent state), odd count, then a plain call
    assert up.prove_batch(p, variants[:3]) == want[:3]
    assert p.prove(variants[4]) == want[4]
    assert p.prove_resident() == want[4]
    assert up.prove_batch(p, []) == []
    p.close()
