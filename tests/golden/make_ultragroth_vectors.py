"""Generates tests/golden/ultragroth_ref_vectors.json from the REFERENCE's own code (needs /root/reference; run here, commit the output):

    python -m tests.golden.make_ultragroth_vectors

The reference holds no UltraGroth fixture, so these vectors pin the path to reference-held CODE instead (oracle/_ref/libref_host.so,
built by oracle/ref_build.py from the sources under /root/reference, entry points in oracle/ref_wrap_host.cpp):
  * the synthetic protocol-1337 zkey and .uwtns (and a protocol-1 zkey) parsed by ZKeyUtils::ultra_groth_loadHeader / loadHeader
    (src/zkey_utils.cpp:42-76,123-163) and WtnsUtils::loadHeader (src/wtns_utils.cpp:13-26) over BinFileUtils::BinFile: every
    header field, header-point offsets and section extents as the reference sees them;
  * compute_lookup (src/ultra_groth.cpp:62-106) digits from the reference's RawFr -- RawFr::set(int), mpz_invert-based inv --
    including frequencies >= 2^31 (negative after the int conversion) and a frequency of 0;
  * derive_challenge (src/ultra_groth.cpp:32-58) from RawFq::toMpz, FIPS202_KECCAK_256, RawFr::fromMpz;
  * RawFr::set(int) and RawFr::inv on single values.
The inputs are regenerated deterministically by the tests (oracle/cpu_synth.cpp with the recorded arguments; sha256 recorded)."""
import hashlib
import json
import os
import random

from oracle import bn254 as bn
from oracle import cpu_prover as cp
from oracle import formats as fm
from oracle import ref_host as rh

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "ultragroth_ref_vectors.json")
ULTRA_ARGS = dict(logn=7, seed=33, ultra=True, lookup_bits=3, n_chunks=20)
G16_ARGS = dict(logn=5, seed=3, ultra=False)


def u32s(b):
    return [int.from_bytes(b[i:i + 4], "little") for i in range(0, len(b), 4)]


def lookup_case(wt_bytes, rand_mont, patch_freqs=None):
    wf = fm.BinFile(wt_bytes, "wtns", 2)
    signals = bytes(wf.section(2))
    chunks, freqs, widx, pidx = (u32s(bytes(wf.section(i))) for i in (3, 4, 5, 6))
    if patch_freqs:
        for i, v in patch_freqs.items():
            freqs[i] = v
    filled = rh.compute_lookup(signals, chunks, freqs, widx, pidx, rand_mont)
    return {"rand_mont": hex(rand_mont), "freqs": freqs, "filled_sha256": hashlib.sha256(filled).hexdigest(),
            "written": {str(i): filled[32 * i:32 * i + 32].hex() for i in sorted(set(widx))}}


def main():
    assert rh.available(), "oracle/_ref/libref_host.so could not be built (is /root/reference here?)"
    rng = random.Random(20261105)
    zk, wt, _ = cp.synth(threads=2, **ULTRA_ARGS)
    zk16, wt16, _ = cp.synth(threads=2, **G16_ARGS)
    out = {
        "generator": "tests/golden/make_ultragroth_vectors.py over oracle/_ref/libref_host.so (reference code compiled in place)",
        "ultra_args": ULTRA_ARGS, "groth16_args": G16_ARGS,
        "ultra_zkey_sha256": hashlib.sha256(zk.tobytes()).hexdigest(), "ultra_wtns_sha256": hashlib.sha256(wt.tobytes()).hexdigest(),
        "groth16_zkey_sha256": hashlib.sha256(zk16.tobytes()).hexdigest(),
        "ultra_zkey_header": rh.zkey_header(zk, True), "ultra_wtns_header": rh.wtns_header(wt),
        "groth16_zkey_header": rh.zkey_header(zk16, False), "groth16_wtns_header": rh.wtns_header(wt16),
    }
    r1 = rng.randrange(bn.R) * bn.MONT_R % bn.R
    r2 = rng.randrange(bn.R) * bn.MONT_R % bn.R
    out["compute_lookup"] = [
        lookup_case(wt.tobytes(), r1),
        lookup_case(wt.tobytes(), r2, {0: 0x80000000, 1: 0xfffffffe, 2: 0, 3: 0x7fffffff, 5: 0x80000005}),
    ]
    ch = []
    # k = 1 is left out on purpose: its x = 1 needs ONE 64-bit word and mpz_export (src/ultra_groth.cpp:43,46) writes only the words a
    # value needs, left-aligned, leaving the rest of the reference's stack buffer uninitialised -- undefined for x or y < 2^192
    # (probability 2^-62 for a real commitment); every restatement here writes the full 32 big-endian bytes
    for k in (2, 3, 0xB200, rng.randrange(bn.R)):
        P = bn.G1.mul_affine(bn.G1_GEN, k)
        xy = (P[0] * bn.MONT_R % bn.Q).to_bytes(32, "little") + (P[1] * bn.MONT_R % bn.Q).to_bytes(32, "little")
        ch.append({"k": hex(k), "commitment_xy_mont": xy.hex(), "rand_mont": hex(rh.derive_challenge(xy))})
    out["derive_challenge"] = ch
    out["fr_set_int"] = {str(v): hex(rh.fr_set_int(v)) for v in (-5, -1, 0, 1, 65536, 2**31 - 1, -2**31)}
    out["fr_inv"] = []
    for _ in range(4):
        a = rng.randrange(1, bn.R)
        out["fr_inv"].append({"a_mont": hex(a), "inv_mont": hex(rh.fr_inv(a))})
    with open(OUT, "w") as fh:
        json.dump(out, fh, indent=1)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
