"""CPU tests (no GPU): the oracle is pinned against everything the reference offers for this path.

  * golden vectors produced by the reference's OWN field code and keccak (tests/golden/field_vectors.json,
    made by tests/golden/make_field_vectors.py from build/f{r,q}_raw_generic.cpp + src/keccak256.cpp)
  * the reference's only fixture, testdata/ (committed under tests/golden/ref_testdata): header points decode
    to verification_key.json; prove -> verify -> tampered public input must fail (the reference's CI job,
    .github/workflows/build.yml:69-81)
  * the C++ restatement (oracle/cpu_prover.cpp, also the timed CPU baseline) against the Python big-int oracle
  * the device limb algorithms (ff.cuh / curve.cuh) compiled for the host with a software carry flag
"""
import ctypes
import json
import os
import random
import subprocess

import numpy as np
import pytest

from oracle import bn254 as bn
from oracle import cpu_prover as cp
from oracle import groth16 as og
from oracle.formats import ZKey, Wtns
from tests.util import (chain_points, fe_list_bytes, g1_bytes, g1_from_bytes, g2_bytes, g2_from_bytes, scalars_uniform,
                        scalars_witness_like)

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
OPS = ["mul", "add", "sub", "neg", "to_mont", "from_mont"]


@pytest.fixture(scope="module")
def vectors():
    return json.load(open(os.path.join(GOLD, "field_vectors.json")))


@pytest.fixture(scope="module")
def testdata():
    zk = open(os.path.join(GOLD, "ref_testdata", "circuit_final.zkey"), "rb").read()
    wt = open(os.path.join(GOLD, "ref_testdata", "witness.wtns"), "rb").read()
    vk = og.vk_from_snarkjs_json(open(os.path.join(GOLD, "ref_testdata", "verification_key.json")).read())
    return zk, wt, vk


@pytest.fixture(scope="module")
def emul():
    """Host build of the DEVICE limb algorithms (tests/emul/emul.cpp includes ultragroth_b200/csrc/curve.cuh)."""
    out = os.path.join(HERE, "emul", "_build")
    os.makedirs(out, exist_ok=True)
    lib = os.path.join(out, "libemul.so")
    src = os.path.join(HERE, "emul", "emul.cpp")
    deps = [src] + [os.path.join(HERE, "..", "ultragroth_b200", "csrc", f) for f in ("ff.cuh", "curve.cuh", "memops.cuh", "msm.cuh")]
    if not os.path.exists(lib) or any(os.path.getmtime(lib) < os.path.getmtime(d) for d in deps):
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-o", lib, src])
    L = ctypes.CDLL(lib)
    L.emul_field_op.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long]
    L.emul_point_op.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    L.emul_next_bucket.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p]
    L.emul_next_bucket.restype = ctypes.c_uint32
    L.emul_scalar_digits.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    return L


def _ints(hexes):
    return [int(h, 16) for h in hexes]


def test_python_oracle_matches_reference_field_code(vectors):
    for name, mod in (("fr", bn.R), ("fq", bn.Q)):
        rec = vectors["fields"][name]
        xs, ys = _ints(rec["x"]), _ints(rec["y"])
        fns = {"mul": lambda x, y: bn.mont_mul(x, y, mod), "add": lambda x, y: bn.raw_add(x, y, mod),
               "sub": lambda x, y: bn.raw_sub(x, y, mod), "neg": lambda x, y: bn.raw_neg(x, mod),
               "to_mont": lambda x, y: bn.to_mont(x, mod), "from_mont": lambda x, y: bn.from_mont(x, mod)}
        for op in OPS:
            assert _ints(rec[op]) == [fns[op](x, y) for x, y in zip(xs, ys)], (name, op)
        assert _ints(rec["unreduced_mul"]) == [bn.mont_mul(x, y, mod) for x, y in zip(_ints(rec["unreduced_x"]), _ints(rec["unreduced_y"]))]


def test_cpp_oracle_and_device_limbs_match_reference_field_code(vectors, emul):
    for fid, (name, mod) in enumerate((("fr", bn.R), ("fq", bn.Q))):
        rec = vectors["fields"][name]
        for xs, ys, want in ((_ints(rec["x"]), _ints(rec["y"]), _ints(rec["mul"])),
                             (_ints(rec["unreduced_x"]), _ints(rec["unreduced_y"]), _ints(rec["unreduced_mul"]))):
            a, b = fe_list_bytes(xs), fe_list_bytes(ys)
            got = cp.field_mul(fid, a, b)
            assert [int.from_bytes(got[i:i + 32], "little") for i in range(0, len(got), 32)] == want
        xs, ys = _ints(rec["x"]), _ints(rec["y"])
        a, b = fe_list_bytes(xs), fe_list_bytes(ys)
        for op, opname in enumerate(OPS):
            out = ctypes.create_string_buffer(len(a))
            assert emul.emul_field_op(fid, op, a, b, out, len(xs)) == 0
            assert [int.from_bytes(out.raw[i:i + 32], "little") for i in range(0, len(a), 32)] == _ints(rec[opname]), (name, opname)
        # inversion (device fp_inv) against pow
        nz = [x for x in xs if x % mod][:20]
        out = ctypes.create_string_buffer(32 * len(nz))
        mont = [x * bn.MONT_R % mod for x in nz]
        emul.emul_field_op(fid, 7, fe_list_bytes(mont), fe_list_bytes(mont), out, len(nz))
        assert [int.from_bytes(out.raw[i:i + 32], "little") for i in range(0, 32 * len(nz), 32)] == [pow(x, -1, mod) * bn.MONT_R % mod for x in nz]


def test_keccak_matches_reference(vectors):
    for k in vectors["keccak256"]:
        msg = bytes.fromhex(k["msg"])
        assert bn.keccak256(msg).hex() == k["digest"]
        assert cp.keccak256(msg).hex() == k["digest"]


def _xyzz_bytes(P, group):
    """affine oracle point -> XYZZ bytes (ZZ = ZZZ = 1) in the device layout"""
    one = bn.MONT_R % bn.Q
    if group == 1:
        if P is None:
            return b"\0" * 128
        return g1_bytes(P) + one.to_bytes(32, "little") * 2
    if P is None:
        return b"\0" * 256
    return g2_bytes(P) + (one.to_bytes(32, "little") + b"\0" * 32) * 2


def _xyzz_to_affine(b, group):
    F = bn.G1.F if group == 1 else bn.G2.F
    n = 32 if group == 1 else 64
    rd = (lambda s: int.from_bytes(s, "little") * bn.FQ_RINV % bn.Q) if group == 1 else \
        (lambda s: (int.from_bytes(s[:32], "little") * bn.FQ_RINV % bn.Q, int.from_bytes(s[32:], "little") * bn.FQ_RINV % bn.Q))
    X, Y, ZZ, ZZZ = (rd(b[i * n:(i + 1) * n]) for i in range(4))
    if F.is_zero(ZZ):
        return None
    return (F.mul(X, F.inv(ZZ)), F.mul(Y, F.inv(ZZZ)))


@pytest.mark.parametrize("group", [1, 2])
def test_device_curve_formulas_on_host(emul, group):
    """madd / add / dbl / to_affine of curve.cuh, including the doubling, cancellation and infinity branches."""
    C, gen = (bn.G1, bn.G1_GEN) if group == 1 else (bn.G2, bn.G2_GEN)
    enc = g1_bytes if group == 1 else g2_bytes
    dec = g1_from_bytes if group == 1 else g2_from_bytes
    sz = 128 if group == 1 else 256
    P, Q2 = C.mul_affine(gen, 5), C.mul_affine(gen, 11)
    negP = (P[0], C.F.neg(P[1]))

    def run(op, acc, q, neg=0):
        out = ctypes.create_string_buffer(sz)
        emul.emul_point_op(group, op, acc, q, neg, out)
        return out.raw

    accP = _xyzz_bytes(P, group)
    assert _xyzz_to_affine(run(0, accP, enc(Q2)), group) == C.mul_affine(gen, 16)
    assert _xyzz_to_affine(run(0, accP, enc(Q2), 1), group) == C.to_affine(C.add(C.to_jac(P), C.neg(C.to_jac(Q2))))
    assert _xyzz_to_affine(run(0, accP, enc(P)), group) == C.mul_affine(gen, 10)       # P + P -> doubling branch
    assert _xyzz_to_affine(run(0, accP, enc(negP)), group) is None                     # P - P -> infinity
    assert _xyzz_to_affine(run(0, accP, enc(None)), group) == P                        # + infinity
    assert _xyzz_to_affine(run(0, _xyzz_bytes(None, group), enc(Q2)), group) == Q2     # infinity + Q
    s16 = run(0, accP, enc(Q2))
    assert _xyzz_to_affine(run(1, s16, accP), group) == C.mul_affine(gen, 21)          # general add
    assert _xyzz_to_affine(run(1, s16, s16), group) == C.mul_affine(gen, 32)           # add -> doubling
    assert _xyzz_to_affine(run(2, s16, s16), group) == C.mul_affine(gen, 32)           # dbl
    aff = ctypes.create_string_buffer(sz // 2)
    emul.emul_point_op(group, 3, s16, s16, 0, aff)
    assert dec(aff.raw) == C.mul_affine(gen, 16)


def test_testdata_header_matches_verification_key(testdata):
    zk_bytes, wt_bytes, vk = testdata
    zk = ZKey(zk_bytes)
    assert (zk.protocol, zk.nVars, zk.nPublic, zk.domainSize) == (1, 1003, 1, 1024)
    assert zk.alpha1 == vk["alpha1"] and zk.beta2 == vk["beta2"] and zk.delta2 == vk["delta2"]
    assert zk.ic() == vk["IC"]
    wt = Wtns(wt_bytes)
    assert wt.nVars == 1003 and wt.prime == bn.R and wt.signals()[0] == 1


def test_reference_ci_flow_prove_verify_tamper(testdata):
    """.github/workflows/build.yml:69-81 on the oracle: prove, verify, decrement public[0] -> must fail."""
    zk, wt, vk = testdata
    proof, public = og.groth16_prove(zk, wt, 0xabcdef0123456789, 0x123456789abcdef)
    assert og.groth16_verify(vk, proof, public)
    assert not og.groth16_verify(vk, proof, [(public[0] - 1) % bn.R])
    js = og.proof_to_json(proof)
    assert og.proof_from_json(js)["pi_a"] == proof["pi_a"] and len(js) < 810


def test_cpp_oracle_matches_python_oracle_on_testdata(testdata):
    zk, wt, vk = testdata
    r, s = 31337, (1 << 247) + 5
    proof, public, inter = og.groth16_prove(zk, wt, r, s, want_intermediates=True)
    res = cp.prove(zk, wt, r, s, want_h=True, domain_size=1024, threads=2)
    p = res["points"]
    assert g1_from_bytes(p[0:64]) == proof["pi_a"]
    assert g2_from_bytes(p[64:192]) == proof["pi_b"]
    assert g1_from_bytes(p[192:256]) == proof["pi_c"]
    assert [int.from_bytes(res["h"][i:i + 32], "little") for i in range(0, 32 * 1024, 32)] == inter["h"]


def _check_msms(msms, inter, ultra):
    assert g1_from_bytes(msms["a"]) == inter["msm_a"]
    assert g1_from_bytes(msms["b1"]) == inter["msm_b1"]
    assert g2_from_bytes(msms["b2"]) == inter["msm_b2"]
    assert g1_from_bytes(msms["c"]) == inter["msm_c"]
    assert g1_from_bytes(msms["h"]) == inter["msm_h"]
    if ultra:
        assert g1_from_bytes(msms["round_c"]) == inter["msm_round"]


def test_cpp_oracle_raw_msm_results_on_testdata(testdata):
    """The per-MSM taps the full-size GPU parity tests compare against (cpu_prove2's out_msms) equal the big-int oracle's."""
    zk, wt, vk = testdata
    r, s = 99991, (1 << 200) + 3
    _, _, inter = og.groth16_prove(zk, wt, r, s, want_intermediates=True)
    res = cp.prove(zk, wt, r, s, threads=2, want_msms=True)
    _check_msms(res["msms"], inter, False)


@pytest.mark.parametrize("ultra,logn", [(False, 6), (True, 7), (True, 9)])
def test_cpu_synth_setup_is_valid_and_cpp_oracle_matches_python(ultra, logn):
    """oracle/cpu_synth.cpp (the CPU twin of the product's synthetic-setup generator): the key is a VALID setup -- the oracle's
    proof verifies under the reference's verifier equation (src/groth16.cpp:298-364, src/ultra_groth.cpp:581-648) and a tampered
    public input fails -- and on it the C++ oracle equals the Python big-int oracle: raw MSMs, H scalars, the witness after
    compute_lookup (src/ultra_groth.cpp:62-106), proof points."""
    zk, wt, vkj = cp.synth(logn, seed=900 + logn, ultra=ultra, lookup_bits=3, n_chunks=40, threads=2)
    zk, wt = zk.tobytes(), wt.tobytes()
    rk, r, s = (1 << 199) + 7, 424242, (1 << 247) + 1
    n_vars = int.from_bytes(wt[60:64], "little")
    if ultra:
        proof, public, inter = og.ultragroth_prove(zk, wt, rk, r, s, want_intermediates=True)
        assert og.ultragroth_verify(og.vk_from_ultragroth_json(vkj), proof, public)
        assert not og.ultragroth_verify(og.vk_from_ultragroth_json(vkj), proof, [(public[0] + 1) % bn.R])
    else:
        proof, public, inter = og.groth16_prove(zk, wt, r, s, want_intermediates=True)
        assert og.groth16_verify(og.vk_from_snarkjs_json(vkj), proof, public)
        assert not og.groth16_verify(og.vk_from_snarkjs_json(vkj), proof, [(public[0] + 1) % bn.R])
    res = cp.prove(zk, wt, r, s, rk, threads=2, want_h=True, want_witness=True, domain_size=1 << logn, n_vars=n_vars, want_msms=True)
    _check_msms(res["msms"], inter, ultra)
    assert [int.from_bytes(res["h"][i:i + 32], "little") for i in range(0, 32 << logn, 32)] == inter["h"]
    p = res["points"]
    assert g1_from_bytes(p[0:64]) == proof["pi_a"] and g2_from_bytes(p[64:192]) == proof["pi_b"]
    assert g1_from_bytes(p[192:256]) == proof["pi_f" if ultra else "pi_c"]
    if ultra:
        assert g1_from_bytes(p[256:320]) == proof["pi_r"]
        assert [int.from_bytes(res["witness"][i:i + 32], "little") for i in range(0, 32 * n_vars, 32)] == [x % bn.R for x in inter["witness"]]


def test_lookup_frequency_int_conversion_in_both_oracles():
    """RawFr::field.mul(frequencies[i], inv) converts the uint32_t frequency to `int` (build/fr.hpp:251 -> RawFr::set(int),
    build/fr.cpp:209-223): a frequency >= 2^31 is a negative number.  Both oracles restate that."""
    from oracle import formats as fm
    zk, wt, _ = cp.synth(7, seed=33, ultra=True, lookup_bits=3, n_chunks=20, threads=2)
    zk, wt = zk.tobytes(), bytearray(wt.tobytes())
    off4 = fm.BinFile(bytes(wt), "wtns", 2).sections[4][0][0]
    wt[off4 + 4:off4 + 8] = (0xfffffffe).to_bytes(4, "little")  # frequency[1] = -2
    wt = bytes(wt)
    n_vars = int.from_bytes(wt[60:64], "little")
    _, _, inter = og.ultragroth_prove(zk, wt, 5, 6, 7, want_intermediates=True)
    res = cp.prove(zk, wt, 6, 7, 5, threads=2, want_witness=True, n_vars=n_vars)
    got = [int.from_bytes(res["witness"][i:i + 32], "little") for i in range(0, 32 * n_vars, 32)]
    assert got == [x % bn.R for x in inter["witness"]]
    # prod_1 = -2 / (1 + rand): the last T signals are the prod values (synthetic layout), T = 8
    rand = inter["rand"] % bn.R
    assert got[n_vars - 8 + 1] == (bn.R - 2) * pow((1 + rand) % bn.R, -1, bn.R) % bn.R


@pytest.mark.parametrize("group", [1, 2])
def test_cpp_msm_vs_python(group):
    C, gen = (bn.G1, bn.G1_GEN) if group == 1 else (bn.G2, bn.G2_GEN)
    enc = g1_bytes if group == 1 else g2_bytes
    dec = g1_from_bytes if group == 1 else g2_from_bytes
    for n, sc_fn in ((1, scalars_uniform), (50, scalars_uniform), (1500, scalars_witness_like)):
        pts = chain_points(C, gen, n, seed=n)
        pts[n // 2] = None if n > 1 else pts[0]
        sc = sc_fn(n, seed=n + 1)
        got = dec(cp.msm(group, b"".join(enc(P) for P in pts), fe_list_bytes(sc), n, threads=2))
        assert got == C.to_affine(C.msm(pts, sc))


@pytest.mark.parametrize("logn", [0, 1, 3, 8, 11])
def test_cpp_ntt_vs_python(logn):
    n = 1 << logn
    rng = random.Random(logn)
    xs = [rng.randrange(bn.R) for _ in range(n)]
    mont = fe_list_bytes([x * bn.MONT_R % bn.R for x in xs])
    un = lambda b: [int.from_bytes(b[i:i + 32], "little") * bn.FR_RINV % bn.R for i in range(0, len(b), 32)]
    assert un(cp.ntt(mont, logn, threads=2)) == bn.ntt(xs)
    assert un(cp.ntt(mont, logn, inverse=True, threads=2)) == bn.ntt(xs, inverse=True)


def test_ntt_root_convention():
    """omega_{2^28} = 5^((r-1)/2^28) (ffiasm / snarkjs convention); a wrong root makes the testdata proof fail."""
    w = bn.fr_root(28)
    assert pow(w, 1 << 28, bn.R) == 1 and pow(w, 1 << 27, bn.R) == bn.R - 1
    assert bn.fr_root(3) == pow(w, 1 << 25, bn.R)


def test_oracle_reproduces_golden_testdata_proof():
    """tests/golden/testdata_proof.json freezes the proof of the reference fixture for fixed r, s (generator:
    tests/golden/make_proof_vectors.py): the oracle must keep producing exactly these bytes, and the product's verifier
    must accept them against the reference's verification_key.json."""
    import hashlib
    import json
    import os
    from oracle import groth16 as og
    from ultragroth_b200 import verifier as uv
    gold = os.path.join(os.path.dirname(__file__), "golden")
    g = json.load(open(os.path.join(gold, "testdata_proof.json")))
    zk = open(os.path.join(gold, "ref_testdata", "circuit_final.zkey"), "rb").read()
    wt = open(os.path.join(gold, "ref_testdata", "witness.wtns"), "rb").read()
    proof, public, inter = og.groth16_prove(zk, wt, int(g["r"], 16), int(g["s"], 16), want_intermediates=True)
    assert json.loads(og.proof_to_json(proof)) == g["proof"]
    assert json.loads(og.public_to_json(public)) == g["public"]
    assert hashlib.sha256(b"".join(int(x).to_bytes(32, "little") for x in inter["h"])).hexdigest() == g["h_sha256"]
    vkj = open(os.path.join(gold, "ref_testdata", "verification_key.json")).read()
    assert uv.groth16_verify(json.dumps(g["proof"]), json.dumps(g["public"]), vkj) == (uv.VERIFIER_VALID_PROOF, "")


def test_karatsuba_product_matches_cios_product_on_cpu(tmp_path):
    """fp_mul_ks (Karatsuba + separate Montgomery reduction, ff_variants.cuh) against fp_mul -- itself pinned to the reference's field code
    above -- on 2*10^5 random and edge-case operand pairs per field, through the host build of the device limb code."""
    import subprocess
    src = tmp_path / "ks.cpp"
    src.write_text(r'''
#include "ff_variants.cuh"
#include <cstdio>
#include <cstdlib>
using namespace ug;
template <class F> int run() {
    srand(7);
    int bad = 0;
    for (int it = 0; it < 200000; it++) {
        F x, y;
        for (int i = 0; i < 8; i++) { x.v[i] = (uint32_t)rand() * 65537u + (uint32_t)rand(); y.v[i] = (uint32_t)rand() * 65537u + (uint32_t)rand(); }
        if (it % 7 == 0) for (int i = 0; i < 8; i++) x.v[i] = 0xffffffffu;
        if (it % 11 == 0) { y = F::modulus(); y.v[0] -= 1; } else { y.v[7] &= 0x1fffffffu; }
        if (it % 13 == 0) for (int i = 0; i < 4; i++) { x.v[i] = x.v[i + 4] = 0xffffffffu; y.v[i] = 0xffffffffu; y.v[4 + i] = (i == 3) ? 0x1fffffffu : 0xffffffffu; }
        if (it % 17 == 0) x = F::zero();
        if (!(fp_mul(x, y) == fp_mul_ks(x, y))) bad++;
    }
    return bad;
}
int main() { int b = run<Fq>() + run<Fr>(); printf("mismatches %d\n", b); return b != 0; }
''')
    exe = tmp_path / "ks"
    root = os.path.dirname(HERE)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(root, "ultragroth_b200", "csrc"), str(src), "-o", str(exe)])
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0 and "mismatches 0" in out.stdout, out.stdout + out.stderr


def test_fp64_assisted_product_matches_cios_product_on_cpu(tmp_path):
    """fp_mul_dfma (ff_variants.cuh: the m*p half of the reduction from exact 52-bit limb products made by round-towards-zero FMAs)
    against fp_mul on the host build of the device limb code (the FMAs run on the host FPU in FE_TOWARDZERO mode): random pairs,
    edge values, and inputs CONSTRUCTED so that x*y mod 2^256 = 2^256 - L with L of every magnitude up to 2^222 -- the case where the
    dropped low columns borrow through bit 256 and the one-unit repair must fire."""
    import subprocess
    src = tmp_path / "dfma.cpp"
    src.write_text(r'''
#include "ff_variants.cuh"
#include <cstdio>
#include <random>
using namespace ug;
static void mulmod256(const uint32_t* a, const uint32_t* b, uint32_t* r) {
    uint32_t out[8];
    unsigned __int128 acc = 0;
    for (int k = 0; k < 8; k++) {
        for (int i = 0; i <= k; i++) acc += (unsigned __int128)a[i] * b[k - i];
        out[k] = (uint32_t)acc;
        acc >>= 32;
    }
    for (int k = 0; k < 8; k++) r[k] = out[k];
}
template <class F> long run(int iters, uint64_t seed) {
    std::mt19937_64 g(seed);
    long bad = 0;
    for (int it = 0; it < iters; it++) {
        F x, y;
        for (int i = 0; i < 8; i += 2) { uint64_t r = g(); x.v[i] = (uint32_t)r; x.v[i + 1] = (uint32_t)(r >> 32); r = g(); y.v[i] = (uint32_t)r; y.v[i + 1] = (uint32_t)(r >> 32); }
        x.v[7] &= 0x3fffffffu; y.v[7] &= 0x3fffffffu; fp_final_sub(x); fp_final_sub(y);
        const int m = it % 9;
        if (m == 0) x = F::zero();
        if (m == 1) { x = F::modulus(); x.v[0] -= 1; y = x; }
        if (m == 2) y = F::one();
        if (m >= 5) {  // x = (2^256 - L) * y^-1 mod 2^256
            y.v[0] |= 1u;
            uint32_t inv[8] = {1, 0, 0, 0, 0, 0, 0, 0};
            for (int rnd = 0; rnd < 9; rnd++) {
                uint32_t t[8], tm[8];
                mulmod256(y.v, inv, t);
                uint64_t borrow = 0;
                for (int k = 0; k < 8; k++) { uint64_t d = (uint64_t)(k == 0 ? 2 : 0) - t[k] - borrow; tm[k] = (uint32_t)d; borrow = (d >> 63) & 1; }
                mulmod256(inv, tm, inv);
            }
            uint32_t L[8], tgt[8];
            const int bits = 1 + (int)(g() % 222);
            for (int k = 0; k < 8; k++) { L[k] = (uint32_t)g(); const int lo = 32 * k; if (lo >= bits) L[k] = 0; else if (lo + 32 > bits) L[k] &= (1u << (bits - lo)) - 1u; }
            if (m == 8) { for (int k = 0; k < 8; k++) L[k] = 0; L[0] = 1; }
            uint64_t borrow = 0;
            for (int k = 0; k < 8; k++) { uint64_t d = (uint64_t)0 - L[k] - borrow; tgt[k] = (uint32_t)d; borrow = (d >> 63) & 1; }
            mulmod256(tgt, inv, x.v);
        }
        if (!(fp_mul(x, y) == fp_mul_dfma(x, y))) bad++;
    }
    return bad;
}
int main() { long b = run<Fq>(60000, 1) + run<Fr>(60000, 2); printf("mismatches %ld\n", b); return b != 0; }
''')
    exe = tmp_path / "dfma"
    root = os.path.dirname(HERE)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-frounding-math", "-I", os.path.join(root, "ultragroth_b200", "csrc"), str(src), "-o", str(exe)])
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0 and "mismatches 0" in out.stdout, out.stdout + out.stderr


def test_lazy_fq2_product_matches_three_full_products_on_cpu(tmp_path):
    """fq2_mul_lazy (curve.cuh: three 512-bit Karatsuba products, two Montgomery reductions -- the product the G2 kernels use) against
    the plain Fq2 product of three complete Fq products, on the host build of the device limb code."""
    import subprocess
    src = tmp_path / "fq2.cpp"
    src.write_text(r'''
#include "curve.cuh"
#include <cstdio>
#include <random>
using namespace ug;
int main() {
    std::mt19937_64 g(5);
    auto rnd = [&](Fq& v) { for (int i = 0; i < 8; i += 2) { uint64_t r = g(); v.v[i] = (uint32_t)r; v.v[i + 1] = (uint32_t)(r >> 32); } v.v[7] &= 0x3fffffffu; fp_final_sub(v); };
    long bad = 0;
    Fq pm1 = Fq::modulus(); pm1.v[0] -= 1;
    for (int it = 0; it < 200000; it++) {
        Fq2 x, y;
        rnd(x.a); rnd(x.b); rnd(y.a); rnd(y.b);
        const int m = it % 11;
        if (m == 0) { x.a = pm1; x.b = pm1; y.a = pm1; y.b = pm1; }
        if (m == 1) x.a = Fq::zero();
        if (m == 2) { x.b = pm1; y.b = pm1; x.a = Fq::zero(); y.a = Fq::zero(); }
        if (m == 3) x = Fq2::zero();
        if (m == 4) y = Fq2::one();
        Fq ac = fp_mul(x.a, y.a), bd = fp_mul(x.b, y.b), s = fp_mul(fp_add(x.a, x.b), fp_add(y.a, y.b));
        Fq2 want, got;
        want.a = fp_sub(ac, bd);
        want.b = fp_sub(fp_sub(s, ac), bd);
        fq2_mul_lazy(x.a, x.b, y.a, y.b, got.a, got.b);
        if (!(want == got)) bad++;
        Fq w2 = fp_sub(fp_mul(x.a, y.a), fp_mul(x.b, y.b));
        if (!(w2 == want.a)) bad++;
    }
    printf("mismatches %ld\n", bad);
    return bad != 0;
}
''')
    exe = tmp_path / "fq2"
    root = os.path.dirname(HERE)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(root, "ultragroth_b200", "csrc"), str(src), "-o", str(exe)])
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0 and "mismatches 0" in out.stdout, out.stdout + out.stderr


def test_next_bucket_equals_the_linear_walk(emul):
    """msm.cuh next_bucket (two probes, then a binary search) against the plain walk it replaced, on bucket offset arrays with long
    runs of EMPTY buckets -- the case that stalled a thread of shard 0 for 2*10^5 steps."""
    import numpy as np
    rng = np.random.default_rng(7)
    for trial in range(40):
        NB = int(rng.integers(3, 5000))
        counts = np.zeros(NB, dtype=np.uint32)
        live = rng.random(NB) < rng.choice([0.001, 0.02, 0.3, 0.9])
        live[int(rng.integers(0, NB))] = True
        counts[live] = rng.integers(1, 9, size=int(live.sum()))
        starts = np.zeros(NB + 1, dtype=np.uint32)
        starts[1:] = np.cumsum(counts)
        M = int(starts[NB])
        end_out = ctypes.c_uint32()
        for _ in range(60):
            pos = int(rng.integers(0, M))
            # any bucket that does not reach beyond pos is a legal starting point (the kernel starts from the bucket that ENDS at pos)
            owner = int(np.searchsorted(starts[1:], pos, side="right"))
            key = int(rng.integers(-1, owner)) if owner > 0 else -1
            want = key + 1
            while starts[want + 1] <= pos:
                want += 1
            got = emul.emul_next_bucket(starts.ctypes.data, NB, key & 0xFFFFFFFF, pos, ctypes.byref(end_out))
            assert got == want == owner and end_out.value == int(starts[want + 1]), (trial, NB, key, pos)


def test_signed_digit_recoding_reconstructs_the_scalar(emul):
    """msm.cuh scalar_reduce + scalar_digits: W = ceil(255 / c) signed digits with |d| <= 2^(c-1) whose weighted sum is the scalar mod r,
    for every window size the planner can pick (and the forced ones), on boundary scalars and random 256-bit values."""
    import numpy as np
    rng = random.Random(11)
    specials = [0, 1, 2, bn.R - 1, bn.R, bn.R + 1, (1 << 256) - 1, (1 << 255), (1 << 254) - 1, (1 << 253), (1 << 128) - 1]
    for c in (4, 8, 13, 15, 16, 17, 19, 20, 22, 23):
        W = -(-255 // c)
        half = 1 << (c - 1)
        specials_c = specials + [half, half + 1, half - 1, (half << c) | half, sum(half << (c * w) for w in range(W)) % (1 << 256),
                                 sum((half + 1) << (c * w) for w in range(W - 1))]
        for s in specials_c + [rng.getrandbits(256) for _ in range(200)]:
            out = (ctypes.c_int32 * W)()
            emul.emul_scalar_digits(int(s).to_bytes(32, "little"), c, W, out)
            digits = list(out)
            assert all(-(half - 1) <= d <= half for d in digits), (c, hex(s), digits)
            assert sum(d << (c * w) for w, d in enumerate(digits)) == s % bn.R, (c, hex(s))
