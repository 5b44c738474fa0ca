"""ugb_msm_g1_noalloc (include/ultragroth_b200.h): the G1 MSM whose scratch is one fixed __device__ array inside the library, for callers
that may not allocate device memory.  It has no counterpart in the reference, so its
oracle is the existing entry point: the affine result is a canonical value, hence the bar is EQUAL BYTES with ugb_msm_g1 on the same
device inputs -- on several calls in a row, because every call shares the scratch array and finds the previous call's contents in it.

This file sorts last on purpose: it passes alone in a fresh process and after the other GPU tests.
Everything goes through the C ABI (Context.msm_g1_noalloc -> ugb_msm_g1_noalloc, Context.msm -> ugb_msm_g1).
"""
import ctypes

import numpy as np
import pytest

from oracle import bn254 as bn
from tests.util import chain_points, fe_list_bytes, g1_bytes, g1_from_bytes, scalars_uniform
from ultragroth_b200._lib import lib

pytestmark = pytest.mark.gpu


def _scalars(n, seed, kind):
    rng = np.random.default_rng(seed)
    b = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    if kind == "uniform":
        b[:, 31] &= 0x1F                     # < 2^253 < r
    elif kind == "witness":                  # many zeros and ones, some small values, the rest full width
        b[:, 31] &= 0x1F
        u = rng.random(n)
        b[u < 0.5] = 0
        one = (u >= 0.5) & (u < 0.7)
        b[one] = 0
        b[one, 0] = 1
        small = (u >= 0.7) & (u < 0.8)
        b[small, 4:] = 0
    elif kind == "any256":                   # raw 256-bit values, most of them >= r: s * P == (s mod r) * P on G1
        pass
    return np.ascontiguousarray(b)


def _bases(ctx, n, seed, holes=()):
    k = _scalars(n, seed, "uniform")
    out = np.empty((n, 64), dtype=np.uint8)
    L = lib()
    assert L.ugb_fixed_base_mul(ctx.handle, 1, k.ctypes.data, n, out.ctypes.data) == 0
    for i in holes:                          # the zkey's point at infinity: all-zero coordinates
        out[i] = 0
    return out


def _both(ctx, bases_h, scalars_h, n):
    bases = ctx.upload(bases_h)
    sc = ctx.upload(scalars_h)
    try:
        return ctx.msm_g1_noalloc(bases, sc, n), ctx.msm(1, bases, sc, n)
    finally:
        bases.free()
        sc.free()


def test_noalloc_msm_equals_msm_on_calls_in_a_row(ctx):
    """Big call, then small calls that reuse a corner of the scratch the big one filled, then big again: equal bytes every time."""
    max_n = int(lib().ugb_msm_noalloc_max_points())
    assert max_n == 16384
    plan = [(5000, "witness", (3, 17, 4999)), (300, "uniform", ()), (max_n, "uniform", (0,)), (2049, "any256", ()), (1, "uniform", ()),
            (5000, "witness", (3, 17, 4999))]
    first = None
    for call, (n, kind, holes) in enumerate(plan):
        bases_h = _bases(ctx, n, seed=1000 + n, holes=holes)
        scalars_h = _scalars(n, seed=2000 + n, kind=kind)
        got, want = _both(ctx, bases_h, scalars_h, n)
        assert got == want, f"call {call}: n = {n} ({kind}) differs from ugb_msm_g1"
        assert any(got), "a random MSM should not be the point at infinity"
        if call == 0:
            first = got
    assert got == first  # the repeated first call, with the other calls' leftovers in the scratch


def test_noalloc_msm_against_the_bigint_oracle(ctx):
    """... and one small case against the Python big-int oracle directly, corner cases included (zero scalars, an infinity base,
    a repeated base: P + P inside one bucket, P - P across the sum)."""
    n = 40
    pts = chain_points(bn.G1, bn.G1_GEN, n, seed=31)
    pts[5] = None
    pts[7] = pts[6]
    pts[9] = (pts[8][0], (bn.Q - pts[8][1]) % bn.Q)   # -P
    sc = scalars_uniform(n, seed=32)
    sc[0] = 0
    sc[7] = sc[6]                      # same point, same scalar: doubling in every window's bucket
    sc[9] = sc[8]                      # k*P + k*(-P) = 0
    sc[11] = bn.R - 1
    bases = ctx.upload(b"".join(g1_bytes(P) for P in pts))
    dsc = ctx.upload(fe_list_bytes(sc))
    got = g1_from_bytes(ctx.msm_g1_noalloc(bases, dsc, n))
    assert got == bn.G1.to_affine(bn.G1.msm(pts, sc))
    bases.free()
    dsc.free()


def test_noalloc_msm_limits(ctx):
    L = lib()
    max_n = int(L.ugb_msm_noalloc_max_points())
    bases = ctx.upload(np.zeros((8, 64), dtype=np.uint8))
    sc = ctx.upload(np.zeros((8, 32), dtype=np.uint8))
    out = ctypes.create_string_buffer(64)
    # beyond the fixed scratch: refused with a message that names the limit (no launch, nothing read)
    assert L.ugb_msm_g1_noalloc(ctx.handle, bases.ptr, sc.ptr, max_n + 1, out) != 0
    msg = L.ugb_last_error().decode()
    assert str(max_n) in msg and "ugb_msm_g1" in msg
    # empty sum and all-infinity bases: the point at infinity (all-zero bytes), like ugb_msm_g1
    assert ctx.msm_g1_noalloc(bases, sc, 0) == b"\0" * 64
    assert ctx.msm_g1_noalloc(bases, sc, 8) == b"\0" * 64 == ctx.msm(1, bases, sc, 8)
    bases.free()
    sc.free()
