"""ctypes binding of libultragroth_b200.so (the C ABI of include/ultragroth_b200.h and include/prover.h).

The library is the product; this module only loads it.  It fails loudly when the shared object is
missing -- there is no Python or CPU implementation to fall back to.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("UGB_LIB") or os.path.join(_HERE, "libultragroth_b200.so")

_lib = None


class UltraGrothError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise UltraGrothError(
                f"{LIB_PATH} is missing: build it with `python -m ultragroth_b200.build` "
                "(nvcc, sm_100a). ultragroth_b200 has no CPU fallback.")
        _lib = ctypes.CDLL(LIB_PATH)
        _declare(_lib)
    return _lib


def _declare(L):
    c = ctypes
    vp, u64, i32 = c.c_void_p, c.c_uint64, c.c_int
    L.ugb_last_error.restype = c.c_char_p
    L.ugb_device_count.restype = i32
    L.ugb_ctx_create.argtypes = [i32, c.POINTER(vp)]
    L.ugb_ctx_destroy.argtypes = [vp]
    L.ugb_ctx_destroy.restype = None
    L.ugb_sync.argtypes = [vp]
    L.ugb_kernel_launches.argtypes = [vp]
    L.ugb_kernel_launches.restype = u64
    L.ugb_stream.argtypes = [vp]
    L.ugb_stream.restype = vp
    L.ugb_malloc.argtypes = [vp, u64, c.POINTER(vp)]
    L.ugb_free.argtypes = [vp, vp]
    L.ugb_memcpy_h2d.argtypes = [vp, vp, vp, u64]
    L.ugb_memcpy_d2h.argtypes = [vp, vp, vp, u64]
    L.ugb_field_vec_op.argtypes = [vp, i32, i32, vp, vp, vp, u64]
    for name in ("ugb_msm_g1", "ugb_msm_g2", "ugb_msm_g1_xyzz", "ugb_msm_g2_xyzz"):
        getattr(L, name).argtypes = [vp, vp, vp, vp, u64, i32, vp]
    L.ugb_msm_g1_noalloc.argtypes = [vp, vp, vp, u64, vp]
    L.ugb_msm_plan.argtypes = [u64, i32, c.POINTER(i32), c.POINTER(i32), c.POINTER(i32)]
    L.ugb_msm_noalloc_max_points.restype = u64
    L.ugb_msm_noalloc_max_points.argtypes = []
    L.ugb_g1_sum_xyzz_to_affine.argtypes = [vp, u64, vp]
    L.ugb_g2_sum_xyzz_to_affine.argtypes = [vp, u64, vp]
    L.ugb_ntt.argtypes = [vp, vp, i32, i32]
    L.ugb_coset_transform.argtypes = [vp, c.POINTER(vp), i32, i32]
    L.ugb_bench_int_pipe.argtypes = [vp, i32, c.POINTER(c.c_double)]
    L.ugb_synth_build.argtypes = [vp, i32, i32, u64, i32, c.c_uint32, c.POINTER(vp), c.POINTER(u64), c.POINTER(vp), c.POINTER(u64), c.POINTER(vp)]
    L.ugb_host_free.argtypes = [vp]
    L.ugb_host_free.restype = None
    L.ugb_fixed_base_mul.argtypes = [vp, i32, vp, u64, vp]
    # prover extensions
    L.ugb_prover_set_blinding.argtypes = [vp, vp, vp, vp]
    L.ugb_prover_last_h.argtypes = [vp, vp, u64]
    L.ugb_prover_last_msms.argtypes = [vp, vp]
    L.ugb_prover_last_witness.argtypes = [vp, vp, u64, u64]
    L.ugb_prover_create_times.argtypes = [vp, c.POINTER(c.c_double)]
    L.ugb_prover_prove_batch.argtypes = [vp, c.c_uint32, vp, vp, vp, vp, vp, vp, vp, c.c_ulonglong]
    L.ugb_prover_inf_points.argtypes = [vp, c.POINTER(u64)]
    L.ugb_prover_last_proof.argtypes = [vp, vp, u64, vp, u64]
    L.ugb_prover_set_timing.argtypes = [vp, i32]
    L.ugb_prover_last_times.argtypes = [vp, c.POINTER(c.c_float), i32]
    L.ugb_prover_kernel_launches.argtypes = [vp]
    L.ugb_prover_kernel_launches.restype = u64
    L.ugb_prover_prove_resident.argtypes = [vp]
    L.ugb_prover_set_profile.argtypes = [vp, i32]
    L.ugb_prover_kernel_stats.argtypes = [vp, vp, u64]
    L.ugb_prover_kernel_stats.restype = u64
    L.ugb_prover_stream.argtypes = [vp]
    L.ugb_prover_stream.restype = vp
    # the reference's prover C API (include/prover.h)
    ull, cp = c.c_ulonglong, c.c_char_p
    pull = c.POINTER(ull)
    for pre in ("groth16", "ultra_groth"):
        getattr(L, pre + "_public_size_for_zkey_buf").argtypes = [vp, ull, pull, vp, ull]
        getattr(L, pre + "_public_size_for_zkey_file").argtypes = [cp, pull, vp, ull]
        getattr(L, pre + "_proof_size").argtypes = [pull]
        getattr(L, pre + "_proof_size").restype = None
        getattr(L, pre + "_prover_create").argtypes = [c.POINTER(vp), vp, ull, vp, ull]
        getattr(L, pre + "_prover_create_zkey_file").argtypes = [c.POINTER(vp), cp, vp, ull]
        getattr(L, pre + "_prover_prove").argtypes = [vp, vp, ull, vp, pull, vp, pull, vp, ull]
        getattr(L, pre + "_prover_destroy").argtypes = [vp]
        getattr(L, pre + "_prover_destroy").restype = None
        getattr(L, pre + "_prover").argtypes = [vp, ull, vp, ull, vp, pull, vp, pull, vp, ull]
        getattr(L, pre + "_prover_zkey_file").argtypes = [cp, vp, ull, vp, pull, vp, pull, vp, ull]


def check(rc):
    if rc != 0:
        raise UltraGrothError(lib().ugb_last_error().decode(errors="replace"))
