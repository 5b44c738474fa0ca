"""UltraGroth pinned to REFERENCE-HELD code (the reference has no UltraGroth fixture).

tests/golden/ultragroth_ref_vectors.json was produced by the reference's own loaders, RawFr / RawFq and keccak compiled in place
(oracle/ref_build.py -> oracle/_ref/libref_host.so; generator tests/golden/make_ultragroth_vectors.py).  Here:
  * CPU (-m "not gpu"): the oracle's parsers, compute_lookup and derive_challenge (Python big-int AND C++) against those vectors,
    and -- when the compiled reference library is present -- against the reference code live;
  * GPU: the CUDA path (k_lookup_inverses / k_lookup_push and the host challenge derivation behind ultra_groth_prover_prove)
    against the same reference code on the witness a real proof leaves in HBM."""
import hashlib
import json
import os

import pytest

from oracle import bn254 as bn
from oracle import cpu_prover as cp
from oracle import formats as fm
from oracle import groth16 as og
from oracle import ref_host as rh

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ultragroth_ref_vectors.json")
R, M = bn.R, bn.MONT_R


@pytest.fixture(scope="module")
def vec():
    return json.load(open(GOLD))


@pytest.fixture(scope="module")
def files(vec):
    zk, wt, _ = cp.synth(threads=2, **vec["ultra_args"])
    assert hashlib.sha256(zk.tobytes()).hexdigest() == vec["ultra_zkey_sha256"], "synthetic generator drifted: regenerate the vectors"
    assert hashlib.sha256(wt.tobytes()).hexdigest() == vec["ultra_wtns_sha256"]
    return zk.tobytes(), wt.tobytes()


def u32s(b):
    return [int.from_bytes(b[i:i + 4], "little") for i in range(0, len(b), 4)]


def test_synthetic_ultragroth_files_parse_under_the_reference_loaders(vec, files):
    """What ZKeyUtils::ultra_groth_loadHeader / WtnsUtils::loadHeader saw (header fields, point offsets, section extents) equals
    what the oracle's parser sees in the same bytes -- so the protocol-1337 layout the generator emits IS the reference's."""
    zk, wt = files
    h = vec["ultra_zkey_header"]
    z = fm.ZKey(zk)
    assert z.protocol == 1337
    assert (h["n8q"], h["n8r"], int(h["qPrime"], 16), int(h["rPrime"], 16)) == (32, 32, bn.Q, bn.R)
    assert (h["nVars"], h["nPublic"], h["domainSize"], h["nCoefs"]) == (z.nVars, z.nPublic, z.domainSize, z.nCoefs)
    assert (h["num_indexes_c1"], h["num_indexes_c2"], h["rand_indx"]) == (z.num_indexes_c1, z.num_indexes_c2, z.rand_indx)
    # header points: the reference's pointers land on the points the oracle decodes
    assert fm.g1_from_zkey(zk, h["alpha1"]) == z.alpha1 and fm.g1_from_zkey(zk, h["beta1"]) == z.beta1
    assert fm.g2_from_zkey(zk, h["beta2"]) == z.beta2 and fm.g2_from_zkey(zk, h["gamma2"]) == z.gamma2
    assert fm.g1_from_zkey(zk, h["round_delta1"]) == z.round_delta1 and fm.g2_from_zkey(zk, h["round_delta2"]) == z.round_delta2
    assert fm.g1_from_zkey(zk, h["final_delta1"]) == z.final_delta1 and fm.g2_from_zkey(zk, h["final_delta2"]) == z.final_delta2
    bf = fm.BinFile(zk, "zkey", 1)
    for sid, (off, size) in h["sections"].items():
        assert [off, size] == list(bf.sections[int(sid)][0])
    # section sizes the prover relies on (src/prover.cpp:237-260)
    sec = {int(k): v[1] for k, v in h["sections"].items()}
    assert sec[5] == sec[6] == 64 * z.nVars and sec[7] == 128 * z.nVars
    assert sec[8] == 64 * z.num_indexes_c1 and sec[9] == 64 * z.num_indexes_c2
    assert sec[10] == 4 * z.num_indexes_c1 and sec[11] == 4 * z.num_indexes_c2 and sec[12] == 64 * z.domainSize
    wh = vec["ultra_wtns_header"]
    w = fm.Wtns(wt)
    assert (wh["n8"], int(wh["prime"], 16), wh["nVars"]) == (32, bn.R, w.nVars)
    wb = fm.BinFile(wt, "wtns", 2)
    for sid, (off, size) in wh["sections"].items():
        assert [off, size] == list(wb.sections[int(sid)][0])


def test_synthetic_groth16_zkey_parses_under_the_reference_loader(vec):
    zk, wt, _ = cp.synth(threads=2, **vec["groth16_args"])
    assert hashlib.sha256(zk.tobytes()).hexdigest() == vec["groth16_zkey_sha256"]
    h = vec["groth16_zkey_header"]
    z = fm.ZKey(zk.tobytes())
    assert (h["nVars"], h["nPublic"], h["domainSize"], h["nCoefs"]) == (z.nVars, z.nPublic, z.domainSize, z.nCoefs)
    assert fm.g1_from_zkey(zk.tobytes(), h["delta1"]) == z.delta1 and fm.g2_from_zkey(zk.tobytes(), h["delta2"]) == z.delta2


def _lookup_inputs(wt, case):
    wf = fm.BinFile(wt, "wtns", 2)
    signals = bytes(wf.section(2))
    chunks, _, widx, pidx = (u32s(bytes(wf.section(i))) for i in (3, 4, 5, 6))
    return signals, chunks, case["freqs"], widx, pidx, int(case["rand_mont"], 16)


def test_compute_lookup_oracles_equal_the_reference_rawfr(vec, files):
    """(i + rand)^-1 and freq * inv digits of RawFr (mpz_invert, RawFr::set(int): frequencies >= 2^31 are negative) -- both oracles."""
    _, wt = files
    for case in vec["compute_lookup"]:
        signals, chunks, freqs, widx, pidx, rand_mont = _lookup_inputs(wt, case)
        got_cpp = cp.compute_lookup(signals, chunks, freqs, widx, pidx, rand_mont)
        assert hashlib.sha256(got_cpp).hexdigest() == case["filled_sha256"]
        for i, hx in case["written"].items():
            assert got_cpp[32 * int(i):32 * int(i) + 32].hex() == hx
        w = [int.from_bytes(signals[i:i + 32], "little") for i in range(0, len(signals), 32)]
        w = og.compute_lookup(w, chunks, freqs, widx, pidx, rand_mont * bn.FR_RINV % R)
        assert b"".join(int(x % R).to_bytes(32, "little") for x in w) == got_cpp
        if rh.available():  # and live against the reference code
            assert rh.compute_lookup(signals, chunks, freqs, widx, pidx, rand_mont) == got_cpp


def test_derive_challenge_oracles_equal_the_reference(vec):
    for c in vec["derive_challenge"]:
        xy = bytes.fromhex(c["commitment_xy_mont"])
        want = int(c["rand_mont"], 16)
        assert cp.derive_challenge(xy) == want
        x = int.from_bytes(xy[:32], "little") * bn.FQ_RINV % bn.Q
        y = int.from_bytes(xy[32:], "little") * bn.FQ_RINV % bn.Q
        assert og.derive_challenge((x, y)) * M % R == want
        if rh.available():
            assert rh.derive_challenge(xy) == want


def test_rawfr_set_int_and_inv_vectors(vec):
    for v, hx in vec["fr_set_int"].items():
        assert int(v) % R * M % R == int(hx, 16)
    for c in vec["fr_inv"]:
        a = int(c["a_mont"], 16) * bn.FR_RINV % R
        assert pow(a, -1, R) * M % R == int(c["inv_mont"], 16)


@pytest.mark.gpu
@pytest.mark.parametrize("patch", [None, {0: 0x80000000, 1: 0xfffffffe, 2: 0, 5: 0x80000005}])
def test_cuda_lookup_fill_and_challenge_equal_the_reference_code(ctx, patch):
    """A real proof through ultra_groth_prover_prove; the round commitment it publishes (pi_r) goes through the REFERENCE's
    derive_challenge, the original witness through the REFERENCE's compute_lookup, and the result must equal the witness the CUDA
    kernels left in HBM (ugb_prover_last_witness) byte for byte.  Needs the prebuilt oracle/_ref/libref_host.so."""
    if not rh.available():
        pytest.skip("oracle/_ref/libref_host.so not shipped")
    from ultragroth_b200 import prover as up
    zk, wt, _ = ctx.synth(9, seed=4321, ultra=True, lookup_bits=5, n_chunks=150)
    zk, wt = zk.tobytes(), bytearray(wt.tobytes())
    wf = fm.BinFile(bytes(wt), "wtns", 2)
    if patch:
        off4 = wf.sections[4][0][0]
        for i, v in patch.items():
            wt[off4 + 4 * i:off4 + 4 * i + 4] = int(v).to_bytes(4, "little")
    wt = bytes(wt)
    wf = fm.BinFile(wt, "wtns", 2)
    signals = bytes(wf.section(2))
    chunks, freqs, widx, pidx = (u32s(bytes(wf.section(i))) for i in (3, 4, 5, 6))
    # the reference's loaders accept the GPU generator's files
    zh = rh.zkey_header(zk, True)
    n_vars = zh["nVars"]
    assert rh.wtns_header(wt)["nVars"] == n_vars
    p = up.Prover(zk, "ultragroth")
    pj, _ = p.prove(wt)
    pi_r = json.loads(pj)["pi_r"]
    xy = (int(pi_r[0]) * bn.MONT_R % bn.Q).to_bytes(32, "little") + (int(pi_r[1]) * bn.MONT_R % bn.Q).to_bytes(32, "little")
    rand_mont = rh.derive_challenge(xy)
    want = rh.compute_lookup(signals, chunks, freqs, widx, pidx, rand_mont)
    assert p.last_witness(0, n_vars) == want
    p.close()
