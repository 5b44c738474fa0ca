"""ctypes binding of oracle/_ref/libref_host.so: the REFERENCE's own file loaders, RawFr / RawFq and keccak, compiled in place by
oracle/ref_build.py (entry points: oracle/ref_wrap_host.cpp).  Test infrastructure: it pins the oracle for UltraGroth, for which
the reference holds no fixture.  The library travels to the GPU box prebuilt; /root/reference is only needed to (re)build it."""
import ctypes
import json
import os

import numpy as np

from . import ref_build

_lib = None


def available():
    return ref_build.build_host() is not None and os.path.exists(ref_build.LIB_HOST)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libref_host.so is missing and /root/reference is not here to build it")
        L = ctypes.CDLL(ref_build.LIB_HOST)
        vp, u64, u32, i32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int
        L.ref_zkey_header_dump.argtypes = [vp, u64, i32, vp, u64]
        L.ref_wtns_header_dump.argtypes = [vp, u64, vp, u64]
        L.ref_derive_challenge.argtypes = [vp, vp]
        L.ref_derive_challenge.restype = None
        L.ref_compute_lookup.argtypes = [vp, vp, u32, vp, u32, vp, vp, u32, vp]
        L.ref_compute_lookup.restype = None
        L.ref_fr_set_int.argtypes = [i32, vp]
        L.ref_fr_set_int.restype = None
        L.ref_fr_inv.argtypes = [vp, vp]
        L.ref_fr_inv.restype = None
        _lib = L
    return _lib


def _buf(b):
    a = np.frombuffer(bytes(b), dtype=np.uint8).copy() if not isinstance(b, np.ndarray) else np.ascontiguousarray(b).view(np.uint8).reshape(-1)
    return a, ctypes.c_void_p(a.ctypes.data)


def zkey_header(zkey, ultra):
    """ZKeyUtils::loadHeader / ultra_groth_loadHeader (src/zkey_utils.cpp:42-76,123-163) over BinFileUtils::BinFile -> dict"""
    a, p = _buf(zkey)
    out = ctypes.create_string_buffer(8192)
    rc = lib().ref_zkey_header_dump(p, a.nbytes, 1 if ultra else 0, out, len(out))
    if rc:
        raise ValueError(out.value.decode(errors="replace"))
    return json.loads(out.value.decode())


def wtns_header(wtns):
    """WtnsUtils::loadHeader (src/wtns_utils.cpp:13-26) -> dict"""
    a, p = _buf(wtns)
    out = ctypes.create_string_buffer(4096)
    rc = lib().ref_wtns_header_dump(p, a.nbytes, out, len(out))
    if rc:
        raise ValueError(out.value.decode(errors="replace"))
    return json.loads(out.value.decode())


def derive_challenge(commitment_xy_mont: bytes) -> int:
    """src/ultra_groth.cpp:32-58 on RawFq::toMpz / FIPS202_KECCAK_256 / RawFr::fromMpz -> the challenge in Montgomery form (integer)"""
    a, p = _buf(commitment_xy_mont)
    out = np.zeros(32, dtype=np.uint8)
    lib().ref_derive_challenge(p, ctypes.c_void_p(out.ctypes.data))
    return int.from_bytes(out.tobytes(), "little")


def compute_lookup(signals: bytes, chunks, freqs, wtns_idx, push_idx, rand_mont: int) -> bytes:
    """src/ultra_groth.cpp:62-106 on the reference's RawFr; signals are raw 32-byte values; returns the filled witness"""
    s, sp = _buf(signals)
    u = lambda v: np.ascontiguousarray(np.asarray(v, dtype=np.uint32))
    c, f, w, q = u(chunks), u(freqs), u(wtns_idx), u(push_idx)
    r, rp = _buf(int(rand_mont).to_bytes(32, "little"))
    ptr = lambda x: ctypes.c_void_p(x.ctypes.data)
    lib().ref_compute_lookup(sp, ptr(c), len(c), ptr(f), len(f), ptr(w), ptr(q), len(w), rp)
    return s.tobytes()


def fr_set_int(v: int) -> int:
    out = np.zeros(32, dtype=np.uint8)
    lib().ref_fr_set_int(int(v), ctypes.c_void_p(out.ctypes.data))
    return int.from_bytes(out.tobytes(), "little")


def fr_inv(a_mont: int) -> int:
    a, p = _buf(int(a_mont).to_bytes(32, "little"))
    out = np.zeros(32, dtype=np.uint8)
    lib().ref_fr_inv(p, ctypes.c_void_p(out.ctypes.data))
    return int.from_bytes(out.tobytes(), "little")
