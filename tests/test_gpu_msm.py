"""GPU parity: G1 / G2 MSM through the C ABI vs the big-int oracle (affine results are canonical,
so bit-exact comparison is algorithm independent)."""
import random

import pytest

from oracle import bn254 as bn
from tests.util import (chain_points, fe_list_bytes, g1_bytes, g1_from_bytes, g2_bytes, g2_from_bytes,
                        scalars_uniform, scalars_witness_like)

pytestmark = pytest.mark.gpu


def _msm(ctx, group, pts, scalars, window_bits=0, index=None):
    enc = g1_bytes if group == 1 else g2_bytes
    dec = g1_from_bytes if group == 1 else g2_from_bytes
    bases = ctx.upload(b"".join(enc(P) for P in pts))
    sc = ctx.upload(fe_list_bytes(scalars))
    idx = None
    if index is not None:
        import numpy as np
        idx = ctx.upload(np.asarray(index, dtype=np.uint32))
    return dec(ctx.msm(group, bases, sc, len(pts), index=idx, window_bits=window_bits))


def _oracle(group, pts, scalars):
    C = bn.G1 if group == 1 else bn.G2
    return C.to_affine(C.msm(pts, scalars))


@pytest.mark.parametrize("group", [1, 2])
@pytest.mark.parametrize("n", [1, 2, 33, 1000])
def test_msm_small_uniform(ctx, group, n):
    C, gen = (bn.G1, bn.G1_GEN) if group == 1 else (bn.G2, bn.G2_GEN)
    pts = chain_points(C, gen, n, seed=100 + n)
    sc = scalars_uniform(n, seed=200 + n)
    assert _msm(ctx, group, pts, sc) == _oracle(group, pts, sc)


@pytest.mark.parametrize("group", [1, 2])
def test_msm_edge_cases(ctx, group):
    C, gen = (bn.G1, bn.G1_GEN) if group == 1 else (bn.G2, bn.G2_GEN)
    pts = chain_points(C, gen, 64, seed=5)
    # empty
    assert _msm(ctx, group, [], []) is None
    # all-zero scalars
    assert _msm(ctx, group, pts, [0] * 64) is None
    # infinity bases (all-zero encoding, as in testdata sections 5/6) are skipped
    pts2 = list(pts)
    pts2[3] = None
    pts2[17] = None
    sc = scalars_uniform(64, seed=6)
    assert _msm(ctx, group, pts2, sc) == _oracle(group, pts2, sc)
    # extreme scalars: r-1, 1, 2^k boundaries, values >= r (reduced mod r like s*P on the r-torsion)
    sc2 = [bn.R - 1, 1, 2, (1 << 253), (1 << 128) - 1, (1 << 16), (1 << 15), (1 << 15) + 1, bn.R - 2] + [0] * 55
    assert _msm(ctx, group, pts, sc2) == _oracle(group, pts, sc2)
    sc3 = [bn.R, bn.R + 5, (1 << 256) - 1] + [7] * 61
    assert _msm(ctx, group, pts, sc3) == _oracle(group, pts, [s % bn.R for s in sc3])
    # repeated base: every point identical -> buckets must handle P+P (doubling) and P-P
    same = [pts[0]] * 64
    sc4 = [1] * 32 + [bn.R - 1] * 32
    assert _msm(ctx, group, same, sc4) is None
    sc5 = [3] * 64
    assert _msm(ctx, group, same, sc5) == C.mul_affine(pts[0], 192)


@pytest.mark.parametrize("window_bits", [4, 9, 13, 16])
def test_msm_window_sizes(ctx, window_bits):
    pts = chain_points(bn.G1, bn.G1_GEN, 500, seed=9)
    sc = scalars_uniform(500, seed=10)
    assert _msm(ctx, 1, pts, sc, window_bits=window_bits) == _oracle(1, pts, sc)


def test_msm_witness_like_skew(ctx):
    """Heavily skewed digits (half zeros, 20% ones): one giant bucket spans many accumulate slices."""
    n = 6000
    pts = chain_points(bn.G1, bn.G1_GEN, n, seed=11)
    sc = scalars_witness_like(n, seed=12)
    assert _msm(ctx, 1, pts, sc) == _oracle(1, pts, sc)
    ones = [1] * n
    assert _msm(ctx, 1, pts, ones) == _oracle(1, pts, ones)


def test_msm_indexed_scalars(ctx):
    """Gathered scalars (src/ultra_groth.cpp:415-419): scalar_i = w[index[i]]."""
    rng = random.Random(3)
    n, m = 300, 1000
    pts = chain_points(bn.G1, bn.G1_GEN, n, seed=13)
    w = scalars_uniform(m, seed=14)
    idx = [rng.randrange(m) for _ in range(n)]
    enc_w = ctx.upload(fe_list_bytes(w))
    import numpy as np
    bases = ctx.upload(b"".join(g1_bytes(P) for P in pts))
    got = g1_from_bytes(ctx.msm(1, bases, enc_w, n, index=ctx.upload(np.asarray(idx, dtype=np.uint32))))
    assert got == _oracle(1, pts, [w[i] for i in idx])


def test_msm_linearity_large(ctx):
    """Size-independent property at 2^17 points: bases (a_i * G) with known a_i, so the answer is
    (sum a_i s_i mod r) * G -- one oracle scalar multiplication."""
    n = 1 << 17
    rng = random.Random(21)
    # bases start + i*step  =>  a_i = a0 + i*d
    a0, d = rng.randrange(1, bn.R), rng.randrange(1, bn.R)
    start = bn.G1.mul(bn.G1_GEN, a0)
    step = bn.G1.mul(bn.G1_GEN, d)
    # generate on the fly like chain_points but with known discrete logs
    from tests.util import chain_points as _cp  # noqa: F401
    F = bn.G1.F
    jac, cur = [], start
    for _ in range(n):
        jac.append(cur)
        cur = bn.G1.add(cur, step)
    zs = [P[2] for P in jac]
    pref, acc = [], 1
    for z in zs:
        pref.append(acc)
        acc = acc * z % bn.Q
    inv = pow(acc, -1, bn.Q)
    pts = [None] * n
    for i in range(n - 1, -1, -1):
        zi = inv * pref[i] % bn.Q
        inv = inv * zs[i] % bn.Q
        zi2 = zi * zi % bn.Q
        pts[i] = (jac[i][0] * zi2 % bn.Q, jac[i][1] * zi2 * zi % bn.Q)
    sc = [rng.randrange(bn.R) for _ in range(n)]
    expect_k = sum((a0 + i * d) * s for i, s in enumerate(sc)) % bn.R
    assert _msm(ctx, 1, pts, sc) == bn.G1.mul_affine(bn.G1_GEN, expect_k)


@pytest.mark.parametrize("group,logn", [(1, 22), (2, 20)])
def test_msm_full_size_known_discrete_logs(ctx, group, logn):
    """BASELINE-size parity (2^22 G1 / 2^20 G2 points, uniform full-width scalars): the bases are k_i*G for known k_i (made by
    the GPU fixed-base generator, itself checked against the oracle in test_gpu_synth), so the exact answer is
    (sum k_i*s_i mod r)*G -- one big-int dot product and one oracle scalar multiplication."""
    import ctypes
    import numpy as np
    from ultragroth_b200._lib import check, lib
    n = 1 << logn
    rng = np.random.default_rng(1000 + logn)
    kb = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    kb[:, 31] &= 0x1F
    sb = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    sb[:, 31] &= 0x1F
    bases_h = np.empty((n, 64 if group == 1 else 128), dtype=np.uint8)
    check(lib().ugb_fixed_base_mul(ctx.handle, group, kb.ctypes.data, n, bases_h.ctypes.data))
    bases = ctx.upload(bases_h)
    sc = ctx.upload(sb)
    raw = ctx.msm(group, bases, sc, n)
    got = g1_from_bytes(raw) if group == 1 else g2_from_bytes(raw)
    # dot product over the integers, 64-bit limbs via numpy object arithmetic would be slow: go through Python ints in chunks
    kl = kb.view("<u8").reshape(n, 4)
    sl = sb.view("<u8").reshape(n, 4)
    to_int = lambda a: [int(r[0]) | int(r[1]) << 64 | int(r[2]) << 128 | int(r[3]) << 192 for r in a]
    dot = 0
    for lo in range(0, n, 1 << 16):
        dot += sum(x * y for x, y in zip(to_int(kl[lo:lo + (1 << 16)]), to_int(sl[lo:lo + (1 << 16)])))
    C, gen = (bn.G1, bn.G1_GEN) if group == 1 else (bn.G2, bn.G2_GEN)
    assert got == C.mul_affine(gen, dot % bn.R)
    bases.free()
    sc.free()
