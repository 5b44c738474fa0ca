"""Thin object wrapper over the internal CUDA C ABI (include/ultragroth_b200.h)."""
import ctypes

import numpy as np

from ._lib import check, lib

FIELD_FR, FIELD_FQ = 0, 1
OP_MUL, OP_ADD, OP_SUB, OP_NEG, OP_TO_MONT, OP_FROM_MONT, OP_SQR = range(7)


class DeviceArray:
    def __init__(self, ctx, nbytes):
        self.ctx = ctx
        self.nbytes = int(nbytes)
        p = ctypes.c_void_p()
        check(lib().ugb_malloc(ctx.handle, self.nbytes, ctypes.byref(p)))
        self.ptr = p.value

    def free(self):
        if self.ptr:
            if self.ctx.handle:  # the context may already be closed (arrays kept alive by a traceback): its memory went with it
                lib().ugb_free(self.ctx.handle, self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Context:
    """One CUDA device + stream + workspace (ugb_ctx)."""

    def __init__(self, device=0):
        h = ctypes.c_void_p()
        check(lib().ugb_ctx_create(int(device), ctypes.byref(h)))
        self.handle = h
        self.device = device

    def close(self):
        if self.handle:
            lib().ugb_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # memory ------------------------------------------------------------------------------
    def upload(self, data) -> DeviceArray:
        buf = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8)) if isinstance(data, (bytes, bytearray, memoryview)) \
            else np.ascontiguousarray(data)
        arr = DeviceArray(self, max(buf.nbytes, 16))
        if buf.nbytes:
            check(lib().ugb_memcpy_h2d(self.handle, arr.ptr, buf.ctypes.data, buf.nbytes))
        arr.nbytes = buf.nbytes
        return arr

    def empty(self, nbytes) -> DeviceArray:
        return DeviceArray(self, nbytes)

    def download(self, arr: DeviceArray, nbytes=None) -> bytes:
        n = arr.nbytes if nbytes is None else nbytes
        out = np.empty(n, dtype=np.uint8)
        if n:
            check(lib().ugb_memcpy_d2h(self.handle, out.ctypes.data, arr.ptr, n))
        return out.tobytes()

    def sync(self):
        check(lib().ugb_sync(self.handle))

    @property
    def kernel_launches(self):
        return int(lib().ugb_kernel_launches(self.handle))

    @property
    def stream(self):
        return lib().ugb_stream(self.handle)

    # ops ---------------------------------------------------------------------------------
    def field_vec_op(self, field, op, a: DeviceArray, b: DeviceArray, out: DeviceArray, n):
        check(lib().ugb_field_vec_op(self.handle, field, op, a.ptr, b.ptr if b is not None else None, out.ptr, n))

    def msm(self, group, bases: DeviceArray, scalars: DeviceArray, n, index: DeviceArray = None, window_bits=0,
            xyzz=False, scalar_offset=0) -> bytes:
        size = (64 if group == 1 else 128) * (2 if xyzz else 1)
        out = ctypes.create_string_buffer(size)
        fn = getattr(lib(), f"ugb_msm_g{group}" + ("_xyzz" if xyzz else ""))
        check(fn(self.handle, bases.ptr, scalars.ptr + 32 * scalar_offset, index.ptr if index is not None else None,
                 n, window_bits, out))
        return out.raw

    def msm_g1_noalloc(self, bases: DeviceArray, scalars: DeviceArray, n) -> bytes:
        """ugb_msm_g1_noalloc: the G1 MSM with all of its scratch in a fixed __device__ array (no device allocation on the call);
        n <= lib().ugb_msm_noalloc_max_points().  Not in the reference."""
        out = ctypes.create_string_buffer(64)
        check(lib().ugb_msm_g1_noalloc(self.handle, bases.ptr, scalars.ptr, n, out))
        return out.raw

    def ntt(self, data: DeviceArray, logn, inverse=False):
        """FFT<Fr>::fft / ifft: natural order in and out, in place."""
        check(lib().ugb_ntt(self.handle, data.ptr, logn, 1 if inverse else 0))

    def coset_transform(self, vecs, logn):
        """ifft -> multiply by omega_2n^i -> fft on up to three vectors (src/groth16.cpp:110-140)."""
        arr = (ctypes.c_void_p * len(vecs))(*[v.ptr for v in vecs])
        check(lib().ugb_coset_transform(self.handle, arr, len(vecs), logn))

    def bench_int_pipe(self, kind):
        v = ctypes.c_double()
        check(lib().ugb_bench_int_pipe(self.handle, kind, ctypes.byref(v)))
        return v.value

    def synth(self, logn, seed=0xB200, ultra=False, lookup_bits=4, n_chunks=8):
        """Synthetic circuit + valid trusted setup -> (zkey bytes, wtns bytes, vk json str)."""
        zk, wt, vk = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p()
        zn, wn = ctypes.c_uint64(), ctypes.c_uint64()
        check(lib().ugb_synth_build(self.handle, 1 if ultra else 0, logn, seed, lookup_bits, n_chunks, ctypes.byref(zk), ctypes.byref(zn),
                                    ctypes.byref(wt), ctypes.byref(wn), ctypes.byref(vk)))
        try:
            zkey = np.ctypeslib.as_array(ctypes.cast(zk, ctypes.POINTER(ctypes.c_uint8)), shape=(zn.value,)).copy()
            wtns = np.ctypeslib.as_array(ctypes.cast(wt, ctypes.POINTER(ctypes.c_uint8)), shape=(wn.value,)).copy()
            vkj = ctypes.string_at(vk).decode()
        finally:
            lib().ugb_host_free(zk)
            lib().ugb_host_free(wt)
            lib().ugb_host_free(vk)
        return zkey, wtns, vkj

    def fixed_base_mul(self, group, scalars):
        """[k * G for k in scalars] as zkey-layout bytes."""
        n = len(scalars)
        src = b"".join(int(k).to_bytes(32, "little") for k in scalars)
        out = ctypes.create_string_buffer((64 if group == 1 else 128) * max(n, 1))
        check(lib().ugb_fixed_base_mul(self.handle, group, src, n, out))
        return out.raw[:(64 if group == 1 else 128) * n]
