"""GPU parity at the BASELINE sizes, bit-exact against the oracle's C++ restatement of the reference prover
(oracle/cpu_prover.cpp, itself pinned to the Python big-int oracle and the reference's fixture in tests/test_oracle.py):

    configs[1]  Groth16, 2^20 constraints          src/groth16.cpp:48-203
    configs[2]  UltraGroth, 2^22 constraints       src/ultra_groth.cpp:401-462, with the benchmark's own lookup parameters

With the blinding scalars r, s (and r_k) fixed, every quantity below is a canonical value that does not depend on the algorithm
(an MSM result in affine form, a fully reduced field element), so "equal bytes" is the parity bar the north star names:
    * the five (six) RAW MSM results A, B1, B2, C, H (and UltraGroth's round commitment MSM)   src/groth16.cpp:53-64,153-154
    * all domainSize H-MSM scalars (compared whole, by sha256)                                  src/groth16.cpp:66-148
    * the witness after compute_lookup filled it in                                             src/ultra_groth.cpp:62-106
    * the final proof points (the C API's JSON against the oracle's points)                     src/groth16.cpp:158-250
The proofs go through the reference-facing C API (*_prover_prove, include/prover.h) on the resident-table MSM path the benchmark
times (c = 20 at these sizes, one shared bucket set, multi-block reduction tails); the small-n tests force the largest window, c = 22, with
UGB_MSM_PRE_C.
"""
import hashlib
import json
import os
import random

import numpy as np
import pytest

from oracle import bn254 as bn
from oracle import cpu_prover as cp
from oracle import groth16 as og
from tests.util import bytes_fe_list, fe_list_bytes, g1_from_bytes, g2_from_bytes
from ultragroth_b200 import prover as up

pytestmark = pytest.mark.gpu


def lookup_params(logn):  # bench.py's
    return min(16, max(2, logn - 6)), 1 << max(2, logn - 2)


def _proof_points_from_json(pj, ultra):
    """C API JSON (canonical decimal coordinates) -> the oracle's byte layout A | B | C | R (affine Montgomery)."""
    d = json.loads(pj)
    M, Q = bn.MONT_R, bn.Q
    enc = lambda x: (int(x) * M % Q).to_bytes(32, "little")
    a = enc(d["pi_a"][0]) + enc(d["pi_a"][1])
    b = enc(d["pi_b"][0][0]) + enc(d["pi_b"][0][1]) + enc(d["pi_b"][1][0]) + enc(d["pi_b"][1][1])
    ckey = "pi_f" if ultra else "pi_c"
    c = enc(d[ckey][0]) + enc(d[ckey][1])
    r = (enc(d["pi_r"][0]) + enc(d["pi_r"][1])) if ultra else b"\0" * 64
    return a + b + c + r


def _compare_with_cpu_oracle(zk, wt, protocol, r, s, rk, n, n_vars):
    ultra = protocol == "ultragroth"
    want = cp.prove(zk, wt, r, s, rk, want_h=True, want_witness=True, domain_size=n, n_vars=n_vars, want_msms=True)
    p = up.Prover(zk, protocol)
    p.set_blinding(r, s, rk)
    pj, uj = p.prove(wt)
    got = p.last_msms()
    for name in up.Prover.MSM_NAMES:
        if name == "round_c" and not ultra:
            continue
        assert got[name] == want["msms"][name], f"MSM {name} differs from the CPU oracle"
    assert hashlib.sha256(p.last_h(n)).hexdigest() == hashlib.sha256(want["h"]).hexdigest(), "H scalars differ"
    if ultra:
        assert hashlib.sha256(p.last_witness(0, n_vars)).hexdigest() == hashlib.sha256(want["witness"]).hexdigest(), "filled witness differs"
    assert _proof_points_from_json(pj, ultra) == want["points"], "proof points differ"
    # the device-resident leg (what bench.py's `value` times) must give the very same proof for the same blinding
    pj2, uj2 = p.prove_resident()
    assert (pj2, uj2) == (pj, uj)
    assert p.last_msms() == got
    return p, pj, uj


def test_config1_groth16_2pow20_bit_exact(ctx):
    logn = 20
    zk, wt, vkj = ctx.synth(logn, seed=0xB200 + logn)
    r, s = (1 << 246) + 0xC0FFEE, 0x1234567890abcdef1234567890abcdef1234567890abcdef1234567890ab
    p, pj, uj = _compare_with_cpu_oracle(zk, wt, "groth16", r, s, 0, 1 << logn, 1 << logn)
    public = [int(x) for x in json.loads(uj)]
    assert og.groth16_verify(og.vk_from_snarkjs_json(vkj), og.proof_from_json(pj), public)
    p.close()


def test_config2_ultragroth_2pow22_bit_exact(ctx):
    logn = 22
    lookup_bits, chunks = lookup_params(logn)
    zk, wt, vkj = ctx.synth(logn, seed=0xB200 + logn, ultra=True, lookup_bits=lookup_bits, n_chunks=chunks)
    n_vars = int.from_bytes(wt[60:64].tobytes(), "little")
    rk, r, s = (1 << 200) + 0xB200, (1 << 247) - 19, 0xfeedface << 180
    p, pj, uj = _compare_with_cpu_oracle(zk, wt, "ultragroth", r, s, rk, 1 << logn, n_vars)
    public = [int(x) for x in json.loads(uj)]
    assert og.ultragroth_verify(og.vk_from_ultragroth_json(vkj), og.proof_from_json(pj), public)
    p.close()


@pytest.mark.parametrize("protocol,logn", [("groth16", 10), ("ultragroth", 10)])
def test_forced_big_window_tails_small_n(ctx, protocol, logn, monkeypatch):
    """The planner picks a small window at small n, so the c = 22 branches (2^21 shared buckets, chunked running sums, multi-block
    k_sum_points / k_bit_sums) never run there; UGB_MSM_PRE_C forces them on a circuit the Python big-int oracle proves too."""
    monkeypatch.setenv("UGB_MSM_PRE_C", "22")
    ultra = protocol == "ultragroth"
    zk, wt, _ = ctx.synth(logn, seed=77 + logn, ultra=ultra, lookup_bits=4, n_chunks=200)
    zk, wt = zk.tobytes(), wt.tobytes()
    rk, r, s = (1 << 190) + 5, (1 << 230) + 11, (1 << 247) - 1
    if ultra:
        want_proof, want_public, inter = og.ultragroth_prove(zk, wt, rk, r, s, want_intermediates=True)
    else:
        want_proof, want_public, inter = og.groth16_prove(zk, wt, r, s, want_intermediates=True)
    n_vars = int.from_bytes(wt[60:64], "little")
    _, pj, uj = _compare_with_cpu_oracle(zk, wt, protocol, r, s, rk, 1 << logn, n_vars)
    assert pj == og.proof_to_json(want_proof)
    assert uj == og.public_to_json(want_public)


def test_long_spmv_rows_many_segments(ctx):
    """UltraGroth's 'sum of all lookup terms' constraint is ONE row with a term per looked-up signal: at 2^17 with 30 000 chunks the
    row is cut into 8192-term segments (k_spmv_long, k_spmv_long_finish, k_spmv_fix_c); H scalars and proof against the CPU oracle."""
    logn = 17
    zk, wt, _ = ctx.synth(logn, seed=91, ultra=True, lookup_bits=10, n_chunks=30000)
    n_vars = int.from_bytes(wt[60:64].tobytes(), "little")
    p, _, _ = _compare_with_cpu_oracle(zk, wt, "ultragroth", 12345, 67890, 424242, 1 << logn, n_vars)
    p.close()


def test_lookup_frequency_sign_semantics(ctx):
    """RawFr::field.mul(frequencies[i], inv) converts the uint32_t to `int` (build/fr.hpp:251, build/fr.cpp:209-223), so a frequency
    >= 2^31 is a NEGATIVE field element in the reference.  Patch one frequency of a synthetic .uwtns to 0x80000005 and compare the
    filled witness with both oracles (the proof no longer verifies -- the circuit is violated -- but the prover does not care)."""
    from oracle import formats as fm
    logn = 8
    zk, wt, _ = ctx.synth(logn, seed=17, ultra=True, lookup_bits=3, n_chunks=20)
    zk, wt = zk.tobytes(), bytearray(wt.tobytes())
    off4 = fm.BinFile(bytes(wt), "wtns", 2).sections[4][0][0]
    wt[off4:off4 + 4] = (0x80000005).to_bytes(4, "little")
    wt = bytes(wt)
    n_vars = int.from_bytes(wt[60:64], "little")
    rk, r, s = 99, 1234, 5678
    want = cp.prove(zk, wt, r, s, rk, want_witness=True, n_vars=n_vars)
    _, _, inter = og.ultragroth_prove(zk, wt, rk, r, s, want_intermediates=True)
    p = up.Prover(zk, "ultragroth")
    p.set_blinding(r, s, rk)
    p.prove(wt)
    got = p.last_witness(0, n_vars)
    assert got == want["witness"]
    assert bytes_fe_list(got) == [x % bn.R for x in inter["witness"]]
    p.close()


def test_lookup_index_out_of_range_is_an_error(ctx):
    """The reference trusts the lookup index arrays; here an out-of-range witness index is caught on the device and the C API
    returns PROVER_ERROR instead of writing outside the witness."""
    from oracle import formats as fm
    zk, wt, _ = ctx.synth(8, seed=18, ultra=True, lookup_bits=3, n_chunks=20)
    zk, wt = zk.tobytes(), bytearray(wt.tobytes())
    off5 = fm.BinFile(bytes(wt), "wtns", 2).sections[5][0][0]
    wt[off5 + 8:off5 + 12] = (0x7fffffff).to_bytes(4, "little")
    p = up.Prover(zk, "ultragroth")
    with pytest.raises(up.ProverError) as e:
        p.prove(bytes(wt))
    assert e.value.code == up.PROVER_ERROR and "out of range" in e.value.msg
    p.close()


@pytest.mark.parametrize("logn", [18, 22])
def test_ntt_full_size_vs_cpu_oracle(ctx, logn):
    """Every output of a 2^22 forward and inverse transform against the oracle's radix-2 FFT (oracle/cpu_prover.cpp cpu_ntt)."""
    n = 1 << logn
    rng = np.random.default_rng(1000 + logn)
    raw = rng.integers(0, 1 << 62, size=(n, 4), dtype=np.uint64)  # < 2^254: below r after the top limb is capped
    raw[:, 3] &= np.uint64((1 << 60) - 1)
    data = raw.view(np.uint8).reshape(-1)
    d = ctx.upload(data)
    ctx.ntt(d, logn, inverse=False)
    got = ctx.download(d)
    want = cp.ntt(data, logn, inverse=False)
    assert hashlib.sha256(got).digest() == hashlib.sha256(want).digest()
    ctx.ntt(d, logn, inverse=True)
    assert ctx.download(d) == data.tobytes()
    d2 = ctx.upload(data)
    ctx.ntt(d2, logn, inverse=True)
    assert hashlib.sha256(ctx.download(d2)).digest() == hashlib.sha256(cp.ntt(data, logn, inverse=True)).digest()
    # the fused H-pipeline transform on three vectors (ifft, * omega_2n^i, fft) against the same three oracle steps
    vecs = [data, np.roll(data, 32), np.roll(data, 64)]
    ds = [ctx.upload(v) for v in vecs]
    ctx.coset_transform(ds, logn)
    w2n = bn.fr_root(logn + 1)
    # omega_2n^i in Montgomery form, as bytes, built with the oracle's field multiplication in log2(n) doubling steps
    shift = np.frombuffer(fe_list_bytes([bn.MONT_R % bn.R]), dtype=np.uint8).copy()
    step = w2n
    while shift.nbytes < 32 * n:
        cnt = shift.nbytes // 32
        factor = np.tile(np.frombuffer(fe_list_bytes([pow(w2n, cnt, bn.R) * bn.MONT_R % bn.R]), dtype=np.uint8), cnt)
        shift = np.concatenate([shift, np.frombuffer(cp.field_mul(0, shift, factor), dtype=np.uint8)])
    for v, dv in zip(vecs, ds):
        co = np.frombuffer(cp.ntt(v, logn, inverse=True), dtype=np.uint8)
        co = np.frombuffer(cp.field_mul(0, co, shift), dtype=np.uint8)
        assert hashlib.sha256(ctx.download(dv)).digest() == hashlib.sha256(cp.ntt(co, logn, inverse=False)).digest()
