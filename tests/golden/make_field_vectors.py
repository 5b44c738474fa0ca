"""Generates tests/golden/field_vectors.json from the REFERENCE's own code, compiled in this container by
oracle/ref_build.py (build/fr_raw_generic.cpp, build/fq_raw_generic.cpp, src/keccak256.cpp).
/root/reference does not exist on the GPU box, so the vectors are committed; rerun with
    python tests/golden/make_field_vectors.py
"""
import ctypes
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_build  # noqa: E402

R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
Q = 21888242871839275222246405745257275088696311157297823662689037894645226208583
OPS = ["mul", "add", "sub", "neg", "to_mont", "from_mont"]


def main():
    lib = ctypes.CDLL(ref_build.build())
    lib.ref_field_op.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long]
    lib.ref_keccak256.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
    rng = random.Random(0xB200)
    out = {"source": "rarimo/ultragroth build/f{r,q}_raw_generic.cpp + src/keccak256.cpp compiled by oracle/ref_build.py", "fields": {}}
    for fid, (name, mod) in enumerate((("fr", R), ("fq", Q))):
        edge = [0, 1, 2, mod - 1, mod - 2, (1 << 256) % mod, mod >> 1, (1 << 64) - 1, 1 << 64, (1 << 192) + 5]
        xs = [a for a in edge for _ in edge] + [rng.randrange(mod) for _ in range(40)]
        ys = [b for _ in edge for b in edge] + [rng.randrange(mod) for _ in range(40)]
        # unreduced first operands for mul (src/ultra_groth.cpp:54-55 feeds a raw 256-bit digest)
        ux = [(1 << 256) - 1, mod, mod + 1, 2 * mod + 7] + [rng.randrange(1 << 256) for _ in range(12)]
        uy = [rng.randrange(mod) for _ in ux]
        n = len(xs)
        a = b"".join(x.to_bytes(32, "little") for x in xs)
        b = b"".join(y.to_bytes(32, "little") for y in ys)
        rec = {"x": [hex(x) for x in xs], "y": [hex(y) for y in ys]}
        for op, opname in enumerate(OPS):
            o = ctypes.create_string_buffer(32 * n)
            assert lib.ref_field_op(fid, op, a, b, o, n) == 0
            rec[opname] = [hex(int.from_bytes(o.raw[32 * i:32 * i + 32], "little")) for i in range(n)]
        o = ctypes.create_string_buffer(32 * len(ux))
        lib.ref_field_op(fid, 0, b"".join(x.to_bytes(32, "little") for x in ux), b"".join(y.to_bytes(32, "little") for y in uy), o, len(ux))
        rec["unreduced_x"] = [hex(x) for x in ux]
        rec["unreduced_y"] = [hex(y) for y in uy]
        rec["unreduced_mul"] = [hex(int.from_bytes(o.raw[32 * i:32 * i + 32], "little")) for i in range(len(ux))]
        out["fields"][name] = rec
    msgs = [b"", b"abc", bytes(range(64)), (1).to_bytes(32, "big") + (2).to_bytes(32, "big"), b"\xff" * 135, b"\x01" * 136, b"z" * 300]
    kk = []
    for m in msgs:
        o = ctypes.create_string_buffer(32)
        lib.ref_keccak256(m if m else None, len(m), o)
        kk.append({"msg": m.hex(), "digest": o.raw.hex()})
    out["keccak256"] = kk
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "field_vectors.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=0)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
