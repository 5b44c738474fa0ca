"""Host-side mirror of the reference's prover C API (src/prover.h) over libultragroth_b200.so.

Same names, argument meaning and error behaviour as the C functions; the classes only add the buffer
management a C caller does by hand (src/main_prover.cpp:39-65)."""
import ctypes

from ._lib import lib

PROVER_OK, PROVER_ERROR, PROVER_ERROR_SHORT_BUFFER, PROVER_INVALID_WITNESS_LENGTH = 0, 1, 2, 3
ERR_SIZE = 512

STAGE_NAMES = ("upload", "round", "lookup", "sort_w", "msm_a", "msm_b1", "msm_b2", "msm_c", "spmv", "intt", "ntt",
               "hfinal", "sort_h", "msm_h", "total")


class ProverError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"prover error {code}: {msg}")
        self.code = code
        self.msg = msg


def _buf(data):
    """bytes-like -> (ctypes pointer, size, keepalive)"""
    if isinstance(data, (bytes, bytearray)):
        arr = (ctypes.c_char * len(data)).from_buffer_copy(data) if isinstance(data, bytes) else (ctypes.c_char * len(data)).from_buffer(data)
        return ctypes.cast(arr, ctypes.c_void_p), len(data), arr
    import numpy as np
    a = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
    return ctypes.c_void_p(a.ctypes.data), a.nbytes, a


def _err():
    return ctypes.create_string_buffer(ERR_SIZE)


def _msg(e):
    return e.raw.split(b"\0", 1)[0].decode(errors="replace")


def proof_size(protocol="groth16"):
    v = ctypes.c_ulonglong()
    getattr(lib(), _prefix(protocol) + "_proof_size")(ctypes.byref(v))
    return v.value


def public_size_for_zkey_buf(zkey, protocol="groth16"):
    p, n, keep = _buf(zkey)
    v = ctypes.c_ulonglong()
    e = _err()
    rc = getattr(lib(), _prefix(protocol) + "_public_size_for_zkey_buf")(p, n, ctypes.byref(v), e, ERR_SIZE - 1)
    if rc:
        raise ProverError(rc, _msg(e))
    return v.value


def _prefix(protocol):
    return {"groth16": "groth16", "ultragroth": "ultra_groth", "ultra_groth": "ultra_groth"}[protocol]


class Prover:
    """groth16_prover_create / _prove / _destroy (or the ultra_groth_ twins)."""

    def __init__(self, zkey, protocol="groth16"):
        self.pre = _prefix(protocol)
        self.protocol = protocol
        self.handle = ctypes.c_void_p()
        p, n, keep = _buf(zkey)
        e = _err()
        rc = getattr(lib(), self.pre + "_prover_create")(ctypes.byref(self.handle), p, n, e, ERR_SIZE - 1)
        if rc:
            self.handle = ctypes.c_void_p()
            raise ProverError(rc, _msg(e))
        self.public_size = public_size_for_zkey_buf(zkey, protocol)
        self.proof_size = proof_size(protocol)

    def close(self):
        if self.handle:
            getattr(lib(), self.pre + "_prover_destroy")(self.handle)
            self.handle = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_blinding(self, r=None, s=None, rk=0):
        if r is None or s is None:
            lib().ugb_prover_set_blinding(self.handle, None, None, None)
            return
        enc = lambda x: int(x).to_bytes(32, "little")
        lib().ugb_prover_set_blinding(self.handle, enc(r), enc(s), enc(rk))

    def set_timing(self, on=True):
        lib().ugb_prover_set_timing(self.handle, 1 if on else 0)

    def last_times(self):
        arr = (ctypes.c_float * len(STAGE_NAMES))()
        lib().ugb_prover_last_times(self.handle, arr, len(STAGE_NAMES))
        return dict(zip(STAGE_NAMES, [float(x) for x in arr]))

    def last_h(self, domain_size):
        out = ctypes.create_string_buffer(32 * domain_size)
        rc = lib().ugb_prover_last_h(self.handle, out, 32 * domain_size)
        if rc:
            raise ProverError(rc, "ugb_prover_last_h failed")
        return out.raw

    MSM_NAMES = ("a", "b1", "b2", "c", "h", "round_c")

    def last_msms(self):
        """Raw MSM results of the last proof (before blinding / assembly) as affine Montgomery bytes in zkey layout:
        {a, b1: 64 B; b2: 128 B; c, h, round_c: 64 B} (ugb_prover_last_msms)."""
        out = ctypes.create_string_buffer(448)
        rc = lib().ugb_prover_last_msms(self.handle, out)
        if rc:
            raise ProverError(rc, lib().ugb_last_error().decode(errors="replace"))
        b = out.raw
        return {"a": b[0:64], "b1": b[64:128], "b2": b[128:256], "c": b[256:320], "h": b[320:384], "round_c": b[384:448]}

    def last_witness(self, first, count):
        """`count` signals of the witness as it stands in HBM after the last proof (lookup signals filled in)."""
        out = ctypes.create_string_buffer(32 * count)
        rc = lib().ugb_prover_last_witness(self.handle, out, first, count)
        if rc:
            raise ProverError(rc, lib().ugb_last_error().decode(errors="replace"))
        return out.raw

    def create_times(self):
        arr = (ctypes.c_double * 4)()
        lib().ugb_prover_create_times(self.handle, arr)
        return dict(zip(("upload_ms", "csr_ms", "tables_ms", "total_ms"), [float(x) for x in arr]))

    def inf_points(self):
        arr = (ctypes.c_uint64 * 3)()
        lib().ugb_prover_inf_points(self.handle, arr)
        return {"a": int(arr[0]), "b": int(arr[1]), "c": int(arr[2])}

    def prove_resident(self):
        """One whole proof from the witness (and lookup arrays) already in HBM: no host->device input copy.  -> (proof, public)"""
        rc = lib().ugb_prover_prove_resident(self.handle)
        if rc:
            raise ProverError(rc, lib().ugb_last_error().decode(errors="replace"))
        proof = ctypes.create_string_buffer(self.proof_size + 1)
        public = ctypes.create_string_buffer(self.public_size + 1)
        lib().ugb_prover_last_proof(self.handle, proof, len(proof), public, len(public))
        return proof.value.decode(), public.value.decode()

    def set_profile(self, on=True):
        lib().ugb_prover_set_profile(self.handle, 1 if on else 0)

    def kernel_stats(self):
        """{kernel name: (launches, total ms)} from CUDA events on the prover's stream."""
        n = lib().ugb_prover_kernel_stats(self.handle, None, 0)
        buf = ctypes.create_string_buffer(int(n) + 16)
        lib().ugb_prover_kernel_stats(self.handle, buf, len(buf))
        out = {}
        for line in buf.value.decode().splitlines():
            name, cnt, ms = line.split("\t")
            out[name] = (int(cnt), float(ms))
        return out

    @property
    def kernel_launches(self):
        return int(lib().ugb_prover_kernel_launches(self.handle))

    @property
    def stream(self):
        return lib().ugb_prover_stream(self.handle)

    def prove(self, wtns, proof_capacity=None, public_capacity=None):
        """-> (proof_json: str, public_json: str); raises ProverError with the C API's code."""
        p, n, keep = _buf(wtns)
        pcap = self.proof_size if proof_capacity is None else proof_capacity
        ucap = self.public_size if public_capacity is None else public_capacity
        proof = ctypes.create_string_buffer(pcap + 1)
        public = ctypes.create_string_buffer(ucap + 1)
        ps, us = ctypes.c_ulonglong(pcap), ctypes.c_ulonglong(ucap)
        e = _err()
        rc = getattr(lib(), self.pre + "_prover_prove")(self.handle, p, n, proof, ctypes.byref(ps), public, ctypes.byref(us), e, ERR_SIZE - 1)
        if rc:
            raise ProverError(rc, _msg(e))
        return proof.raw.split(b"\0", 1)[0].decode(), public.raw.split(b"\0", 1)[0].decode()


def prove_batch(prover, wtns_list):
    """ugb_prover_prove_batch: pipelined proofs of several witnesses on one resident prover -> [(proof_json, public_json)]"""
    n = len(wtns_list)
    bufs = [_buf(w) for w in wtns_list]
    wp = (ctypes.c_void_p * n)(*[b[0] for b in bufs])
    ws = (ctypes.c_ulonglong * n)(*[b[1] for b in bufs])
    proofs = [ctypes.create_string_buffer(prover.proof_size + 1) for _ in range(n)]
    publics = [ctypes.create_string_buffer(prover.public_size + 1) for _ in range(n)]
    pp = (ctypes.c_char_p * n)(*[ctypes.cast(p, ctypes.c_char_p) for p in proofs])
    up_ = (ctypes.c_char_p * n)(*[ctypes.cast(p, ctypes.c_char_p) for p in publics])
    ps = (ctypes.c_ulonglong * n)(*([prover.proof_size] * n))
    us = (ctypes.c_ulonglong * n)(*([prover.public_size] * n))
    e = _err()
    rc = lib().ugb_prover_prove_batch(prover.handle, n, wp, ws, pp, ps, up_, us, e, ERR_SIZE - 1)
    if rc:
        raise ProverError(rc, _msg(e))
    return [(p.raw.split(b"\0", 1)[0].decode(), u.raw.split(b"\0", 1)[0].decode()) for p, u in zip(proofs, publics)]


def prove_oneshot(zkey, wtns, protocol="groth16"):
    """groth16_prover / ultra_groth_prover (create + prove + destroy)."""
    pre = _prefix(protocol)
    zp, zn, k1 = _buf(zkey)
    wp, wn, k2 = _buf(wtns)
    pcap = proof_size(protocol)
    ucap = public_size_for_zkey_buf(zkey, protocol)
    proof = ctypes.create_string_buffer(pcap + 1)
    public = ctypes.create_string_buffer(ucap + 1)
    ps, us = ctypes.c_ulonglong(pcap), ctypes.c_ulonglong(ucap)
    e = _err()
    rc = getattr(lib(), pre + "_prover")(zp, zn, wp, wn, proof, ctypes.byref(ps), public, ctypes.byref(us), e, ERR_SIZE - 1)
    if rc:
        raise ProverError(rc, _msg(e))
    return proof.raw.split(b"\0", 1)[0].decode(), public.raw.split(b"\0", 1)[0].decode()
