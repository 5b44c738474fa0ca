"""CPU: the host verifier behind the reference's verifier C API (include/verifier.h; src/verifier.h:22-38) against the
oracle's verifier on (a) the reference's own fixture -- testdata/verification_key.json with a proof made by the oracle
prover from circuit_final.zkey + witness.wtns -- and (b) UltraGroth instances built from known exponents (the reference
holds no UltraGroth fixture).  Mirrors the reference's CI job: prove -> verify -> tamper -> verify fails
(.github/workflows/build.yml:69-81)."""
import json
import os
import random

from oracle import bn254 as bn
from oracle import groth16 as og
from ultragroth_b200 import verifier as uv

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_testdata")


def _fixture():
    zk = open(os.path.join(GOLD, "circuit_final.zkey"), "rb").read()
    wt = open(os.path.join(GOLD, "witness.wtns"), "rb").read()
    vkj = open(os.path.join(GOLD, "verification_key.json")).read()
    proof, public = og.groth16_prove(zk, wt, 12345678901234567890, (1 << 200) + 17)
    return og.proof_to_json(proof), og.public_to_json(public), vkj


def test_groth16_reference_fixture_valid_and_tampered():
    pj, uj, vkj = _fixture()
    assert og.groth16_verify(og.vk_from_snarkjs_json(vkj), og.proof_from_json(pj), [int(x) for x in json.loads(uj)])
    assert uv.groth16_verify(pj, uj, vkj) == (uv.VERIFIER_VALID_PROOF, "")
    # tampered public input
    pub = json.loads(uj)
    bad = json.dumps([str((int(pub[0]) + 1) % bn.R)] + pub[1:])
    assert uv.groth16_verify(pj, bad, vkj)[0] == uv.VERIFIER_INVALID_PROOF
    # tampered proof: pi_a replaced by pi_c (still on the curve)
    d = json.loads(pj)
    d["pi_a"] = d["pi_c"]
    assert uv.groth16_verify(json.dumps(d), uj, vkj)[0] == uv.VERIFIER_INVALID_PROOF
    # a point off the curve
    d = json.loads(pj)
    d["pi_a"][0] = str((int(d["pi_a"][0]) + 1) % bn.Q)
    assert uv.groth16_verify(json.dumps(d), uj, vkj)[0] == uv.VERIFIER_INVALID_PROOF
    # inputs reduce mod r like Fr.fromString
    assert uv.groth16_verify(pj, json.dumps([str(int(pub[0]) + bn.R)] + pub[1:]), vkj)[0] == uv.VERIFIER_VALID_PROOF


def test_groth16_error_paths_match_reference_messages():
    pj, uj, vkj = _fixture()
    assert uv.groth16_verify("{not json", uj, vkj) == (uv.VERIFIER_ERROR, "invalid proof data")
    d = json.loads(pj)
    d["protocol"] = "ultragroth"
    assert uv.groth16_verify(json.dumps(d), uj, vkj) == (uv.VERIFIER_ERROR, "invalid proof data")
    assert uv.groth16_verify(pj, "[]", vkj) == (uv.VERIFIER_ERROR, "invalid inputs data")
    assert uv.groth16_verify(pj, '["12x"]', vkj) == (uv.VERIFIER_ERROR, "invalid inputs data")
    k = json.loads(vkj)
    k["curve"] = "bls12381"
    assert uv.groth16_verify(pj, uj, json.dumps(k)) == (uv.VERIFIER_ERROR, "invalid verification key data")
    k = json.loads(vkj)
    del k["vk_delta_2"]
    assert uv.groth16_verify(pj, uj, json.dumps(k)) == (uv.VERIFIER_ERROR, "invalid verification key data")
    assert uv.groth16_verify(pj, json.dumps(json.loads(uj) + ["5"]), vkj) == (uv.VERIFIER_ERROR, "len(inputs)+1 != len(vk.IC)")
    assert uv.ultra_groth_verify(pj, uj, vkj) == (uv.VERIFIER_ERROR, "invalid proof data")


def _g1s(k):
    return bn.G1.mul_affine(bn.G1_GEN, k % bn.R)


def _g2s(k):
    return bn.G2.mul_affine(bn.G2_GEN, k % bn.R)


def _j1(P):
    return [str(P[0]), str(P[1]), "1"]


def _j2(P):
    return [[str(P[0][0]), str(P[0][1])], [str(P[1][0]), str(P[1][1])], ["1", "0"]]


def _ultragroth_instance(seed, n_inputs):
    """A (proof, inputs, vk) triple that satisfies src/ultra_groth.cpp:581-648 by construction: all points are known
    multiples of the generators, and the final commitment's exponent is solved from the pairing equation."""
    rng = random.Random(seed)
    rnd = lambda: rng.randrange(1, bn.R)
    a, b, alpha, beta, gamma, df, dr, rho, ic_rand = (rnd() for _ in range(9))
    ic = [rnd() for _ in range(n_inputs + 1)]
    inputs = [rnd() for _ in range(n_inputs)]
    Rc = _g1s(rho)
    return rng, (a, b, alpha, beta, gamma, df, dr, rho, ic_rand, ic, inputs, Rc)


def test_ultragroth_constructed_instances():
    for seed, n_inputs in ((1, 1), (2, 3)):
        rng, (a, b, alpha, beta, gamma, df, dr, rho, ic_rand, ic, inputs, Rc) = _ultragroth_instance(seed, n_inputs)
        vk = dict(alpha1=_g1s(alpha), beta2=_g2s(beta), gamma2=_g2s(gamma), final_delta2=_g2s(df), round_delta2=_g2s(dr),
                  IC=[_g1s(k) for k in ic], IC_rand=_g1s(ic_rand))
        # vkx exponent = ic0 + sum in_i*ic_i + rand*ic_rand with the challenge the verifier derives from pi_r
        rand = og.derive_challenge(Rc)
        x = (ic[0] + sum(i * k for i, k in zip(inputs, ic[1:])) + rand * ic_rand) % bn.R
        f = ((a * b - alpha * beta - x * gamma - rho * dr) * pow(df, -1, bn.R)) % bn.R
        proof = dict(pi_a=_g1s(a), pi_b=_g2s(b), pi_f=_g1s(f), pi_r=Rc)
        assert og.ultragroth_verify(vk, proof, inputs), "instance construction is wrong"
        pj = json.dumps(dict(pi_a=_j1(proof["pi_a"]), pi_b=_j2(proof["pi_b"]), pi_f=_j1(proof["pi_f"]), pi_r=_j1(Rc), protocol="ultragroth"))
        vkj = json.dumps(dict(protocol="ultragroth", curve="bn128", nPublic=n_inputs, vk_alpha_1=_j1(vk["alpha1"]), vk_beta_2=_j2(vk["beta2"]),
                              vk_gamma_2=_j2(vk["gamma2"]), vk_delta_c2_2=_j2(vk["final_delta2"]), vk_delta_c1_2=_j2(vk["round_delta2"]),
                              IC=[_j1(P) for P in vk["IC"]], IC_rand=_j1(vk["IC_rand"])))
        uj = json.dumps([str(i) for i in inputs])
        assert uv.ultra_groth_verify(pj, uj, vkj) == (uv.VERIFIER_VALID_PROOF, "")
        bad = json.dumps([str((inputs[0] + 1) % bn.R)] + [str(i) for i in inputs[1:]])
        assert uv.ultra_groth_verify(pj, bad, vkj)[0] == uv.VERIFIER_INVALID_PROOF
        d = json.loads(pj)
        d["pi_r"] = _j1(_g1s(rho + 1))  # changes the challenge AND the pairing
        assert uv.ultra_groth_verify(json.dumps(d), uj, vkj)[0] == uv.VERIFIER_INVALID_PROOF
        assert uv.groth16_verify(pj, uj, vkj) == (uv.VERIFIER_ERROR, "invalid proof data")
