"""ctypes binding of the oracle's C++/OpenMP restatement of the reference prover (oracle/cpu_prover.cpp).
Test infrastructure and the timed CPU baseline; never imported by ultragroth_b200/."""
import ctypes
import os
import time

import numpy as np

from . import build_oracle

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = build_oracle.LIB
        build_oracle.build()  # no-op unless the .so is missing or older than cpu_prover.cpp
        L = ctypes.CDLL(path)
        vp, u64, i32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int
        L.cpu_prove.argtypes = [vp, u64, vp, u64, vp, vp, vp, i32, vp, vp, vp, vp]
        L.cpu_prove2.argtypes = [vp, u64, vp, u64, vp, vp, vp, i32, vp, vp, vp, vp, vp]
        L.cpu_synth_build.argtypes = [i32, i32, u64, i32, ctypes.c_uint32, i32, ctypes.POINTER(vp), ctypes.POINTER(u64), ctypes.POINTER(vp),
                                      ctypes.POINTER(u64), ctypes.POINTER(vp)]
        L.cpu_free.argtypes = [vp]
        L.cpu_free.restype = None
        L.cpu_msm_g1.argtypes = [vp, vp, u64, i32, vp]
        L.cpu_msm_g2.argtypes = [vp, vp, u64, i32, vp]
        L.cpu_ntt.argtypes = [vp, i32, i32, i32]
        L.cpu_field_mul.argtypes = [i32, vp, vp, vp, u64]
        L.cpu_keccak256.argtypes = [vp, u64, vp]
        L.cpu_keccak256.restype = None
        L.cpu_derive_challenge.argtypes = [vp, vp]
        L.cpu_derive_challenge.restype = None
        L.cpu_compute_lookup.argtypes = [vp, vp, u64, vp, u64, vp, vp, u64, vp]
        L.cpu_compute_lookup.restype = None
        _lib = L
    return _lib


def _ptr(a):
    return ctypes.c_void_p(a.ctypes.data)


def _arr(b):
    return np.frombuffer(bytes(b), dtype=np.uint8).copy() if not isinstance(b, np.ndarray) else np.ascontiguousarray(b).view(np.uint8).reshape(-1)


def max_threads():
    return int(lib().cpu_max_threads())


MSM_SLICES = {"a": (0, 64), "b1": (64, 128), "b2": (128, 256), "c": (256, 320), "h": (320, 384), "round_c": (384, 448)}


def prove(zkey, wtns, r, s, rk=0, threads=0, want_h=False, want_witness=False, domain_size=None, n_vars=None, want_msms=False):
    """-> dict(points=bytes(320): A|B|C|R affine Montgomery, h=bytes or None, witness=bytes or None, times_ms=[...],
    msms={a,b1,b2,c,h,round_c: affine Montgomery bytes of the RAW MSM results} or None)"""
    zk, wt = _arr(zkey), _arr(wtns)
    enc = lambda x: np.frombuffer(int(x).to_bytes(32, "little"), dtype=np.uint8).copy()
    rr, ss, kk = enc(r), enc(s), enc(rk)
    pts = np.zeros(320, dtype=np.uint8)
    h = np.zeros(32 * domain_size, dtype=np.uint8) if want_h else None
    w = np.zeros(32 * n_vars, dtype=np.uint8) if want_witness else None
    times = np.zeros(5, dtype=np.float64)
    mm = np.zeros(448, dtype=np.uint8)
    rc = lib().cpu_prove2(_ptr(zk), zk.nbytes, _ptr(wt), wt.nbytes, _ptr(rr), _ptr(ss), _ptr(kk), threads, _ptr(pts),
                          _ptr(h) if want_h else None, _ptr(w) if want_witness else None, _ptr(times), _ptr(mm) if want_msms else None)
    if rc:
        raise RuntimeError(f"cpu_prove failed with code {rc}")
    mb = mm.tobytes()
    return dict(points=pts.tobytes(), h=h.tobytes() if want_h else None, witness=w.tobytes() if want_witness else None,
                times_ms=[float(x) for x in times], msms={k: mb[a:b] for k, (a, b) in MSM_SLICES.items()} if want_msms else None)


def synth(logn, seed=0xB200, ultra=False, lookup_bits=4, n_chunks=8, threads=0):
    """CPU generator of the synthetic circuits + valid setup (oracle/cpu_synth.cpp): the same (zkey, wtns, vk json) as the product's
    ugb_synth_build for the same arguments, without touching the GPU library."""
    zk, wt, vk = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p()
    zn, wn = ctypes.c_uint64(), ctypes.c_uint64()
    rc = lib().cpu_synth_build(1 if ultra else 0, logn, seed, lookup_bits, n_chunks, threads, ctypes.byref(zk), ctypes.byref(zn),
                               ctypes.byref(wt), ctypes.byref(wn), ctypes.byref(vk))
    if rc:
        raise RuntimeError(f"cpu_synth_build failed with code {rc}")
    try:
        zkey = np.ctypeslib.as_array(ctypes.cast(zk, ctypes.POINTER(ctypes.c_uint8)), shape=(zn.value,)).copy()
        wtns = np.ctypeslib.as_array(ctypes.cast(wt, ctypes.POINTER(ctypes.c_uint8)), shape=(wn.value,)).copy()
        vkj = ctypes.string_at(vk).decode()
    finally:
        lib().cpu_free(zk)
        lib().cpu_free(wt)
        lib().cpu_free(vk)
    return zkey, wtns, vkj


def msm(group, bases, scalars, n, threads=0):
    b, s = _arr(bases), _arr(scalars)
    out = np.zeros(64 if group == 1 else 128, dtype=np.uint8)
    (lib().cpu_msm_g1 if group == 1 else lib().cpu_msm_g2)(_ptr(b), _ptr(s), n, threads, _ptr(out))
    return out.tobytes()


def ntt(data, logn, inverse=False, threads=0):
    d = _arr(data).copy()  # cpu_ntt works in place: never on the caller's array
    lib().cpu_ntt(_ptr(d), logn, 1 if inverse else 0, threads)
    return d.tobytes()


def field_mul(field, a, b):
    x, y = _arr(a), _arr(b)
    out = np.zeros_like(x)
    lib().cpu_field_mul(field, _ptr(x), _ptr(y), _ptr(out), x.nbytes // 32)
    return out.tobytes()


def keccak256(data: bytes) -> bytes:
    d = _arr(data) if len(data) else np.zeros(1, dtype=np.uint8)
    out = np.zeros(32, dtype=np.uint8)
    lib().cpu_keccak256(_ptr(d), len(data), _ptr(out))
    return out.tobytes()


def derive_challenge(commitment_xy_mont: bytes) -> int:
    a = _arr(commitment_xy_mont)
    out = np.zeros(32, dtype=np.uint8)
    lib().cpu_derive_challenge(_ptr(a), _ptr(out))
    return int.from_bytes(out.tobytes(), "little")


def compute_lookup(signals: bytes, chunks, freqs, wtns_idx, push_idx, rand_mont: int) -> bytes:
    s = _arr(signals).copy()
    u = lambda v: np.ascontiguousarray(np.asarray(v, dtype=np.uint32))
    c, f, w, q = u(chunks), u(freqs), u(wtns_idx), u(push_idx)
    r = _arr(int(rand_mont).to_bytes(32, "little"))
    lib().cpu_compute_lookup(_ptr(s), _ptr(c), len(c), _ptr(f), len(f), _ptr(w), _ptr(q), len(w), _ptr(r))
    return s.tobytes()


def bench(workload, sample_logn, target_logn, threads=None):
    """Times one CPU proof of the same circuit family at 2^sample_logn constraints (a bounded sample) and scales it
    linearly in the constraint count to the target size.  The synthetic zkey comes from the harness generator."""
    import ultragroth_b200 as ug
    threads = threads or max_threads()
    sample_logn = min(sample_logn, target_logn)
    ctx = ug.Context(int(os.environ.get("UGB_DEVICE", "0")))
    ultra = workload == "ultragroth"
    lookup_bits = min(16, max(2, sample_logn - 6))
    chunks = 1 << max(2, sample_logn - 2)
    zk, wt, _ = ctx.synth(sample_logn, seed=0xB200 + sample_logn, ultra=ultra, lookup_bits=lookup_bits, n_chunks=chunks)
    ctx.close()
    t0 = time.perf_counter()
    res = prove(zk, wt, 12345, 67890, 424242, threads=threads)
    wall = time.perf_counter() - t0
    prove_s = res["times_ms"][3] / 1e3
    scale = float(1 << (target_logn - sample_logn))
    return {"value": 1.0 / (prove_s * scale), "unit": "proofs/s", "cores": threads, "kind": "port",
            "sample": f"one {workload} proof of the same circuit family at 2^{sample_logn} constraints ({prove_s:.2f} s on {threads} threads, "
                      f"wall {wall:.2f} s), scaled linearly x{int(scale)} to 2^{target_logn}",
            "stage_ms": {"msm": res["times_ms"][0], "spmv": res["times_ms"][1], "fft": res["times_ms"][2], "prove": res["times_ms"][3]},
            "note": "C++17/OpenMP restatement of src/groth16.cpp / src/ultra_groth.cpp on generic 64-bit limbs (oracle/cpu_prover.cpp); "
                    "the reference's ffiasm assembly engine is absent from the reference tree and cannot be built here"}
