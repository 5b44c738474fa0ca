#!/usr/bin/env python
"""Generates tests/golden/testdata_proof.json: the Groth16 proof of the reference's own fixture
(testdata/circuit_final.zkey + witness.wtns, copied under tests/golden/ref_testdata) for FIXED blinding scalars r, s,
as computed by the big-int oracle (oracle/groth16.py, which follows src/groth16.cpp:48-203).  The reference holds no
golden proof (its blinding is random, src/groth16.cpp:165-166); this file freezes the oracle's answer so that both the
oracle and the CUDA prover are regression-pinned to the same bytes, and records that the proof verifies against the
reference's testdata/verification_key.json.

    python tests/golden/make_proof_vectors.py
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import groth16 as og  # noqa: E402

R_FIX = 0x1234567890abcdef1234567890abcdef1234567890abcdef1234567890ab
S_FIX = (1 << 247) + 12345


def main():
    gold = os.path.join(HERE, "ref_testdata")
    zk = open(os.path.join(gold, "circuit_final.zkey"), "rb").read()
    wt = open(os.path.join(gold, "witness.wtns"), "rb").read()
    vk = og.vk_from_snarkjs_json(open(os.path.join(gold, "verification_key.json")).read())
    proof, public, inter = og.groth16_prove(zk, wt, R_FIX, S_FIX, want_intermediates=True)
    assert og.groth16_verify(vk, proof, public)
    h = b"".join(int(x).to_bytes(32, "little") for x in inter["h"])
    out = {"r": hex(R_FIX), "s": hex(S_FIX), "proof": json.loads(og.proof_to_json(proof)), "public": json.loads(og.public_to_json(public)),
           "h_sha256": hashlib.sha256(h).hexdigest(), "h_count": len(inter["h"]), "verifies_against": "ref_testdata/verification_key.json"}
    with open(os.path.join(HERE, "testdata_proof.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote testdata_proof.json")


if __name__ == "__main__":
    main()
