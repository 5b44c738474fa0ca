"""GPU parity: Fr/Fq vector kernels vs the oracle's restatement of build/fr_raw_generic.cpp (bit-exact)."""
import random

import pytest

from oracle import bn254 as bn
from tests.util import bytes_fe_list, fe_list_bytes

pytestmark = pytest.mark.gpu


def _run(ctx, field, op, xs, ys):
    a = ctx.upload(fe_list_bytes(xs))
    b = ctx.upload(fe_list_bytes(ys))
    o = ctx.empty(32 * len(xs))
    ctx.field_vec_op(field, op, a, b, o, len(xs))
    return bytes_fe_list(ctx.download(o, 32 * len(xs)))


@pytest.mark.parametrize("field,mod", [(0, bn.R), (1, bn.Q)])
def test_field_ops_bit_exact(ctx, field, mod):
    rng = random.Random(1234 + field)
    edge = [0, 1, 2, mod - 1, mod - 2, (1 << 256) % mod, (1 << 255) % mod, mod >> 1, 0xffffffff, 1 << 32,
            (1 << 64) - 1, (1 << 224) - 1]
    xs, ys = [], []
    for e in edge:
        for e2 in edge:
            xs.append(e)
            ys.append(e2)
    n_rand = 20000
    xs += [rng.randrange(mod) for _ in range(n_rand)]
    ys += [rng.randrange(mod) for _ in range(n_rand)]
    assert _run(ctx, field, 0, xs, ys) == [bn.mont_mul(x, y, mod) for x, y in zip(xs, ys)]
    assert _run(ctx, field, 1, xs, ys) == [bn.raw_add(x, y, mod) for x, y in zip(xs, ys)]
    assert _run(ctx, field, 2, xs, ys) == [bn.raw_sub(x, y, mod) for x, y in zip(xs, ys)]
    assert _run(ctx, field, 3, xs, ys) == [bn.raw_neg(x, mod) for x in xs]
    assert _run(ctx, field, 4, xs, ys) == [bn.to_mont(x, mod) for x in xs]
    assert _run(ctx, field, 5, xs, ys) == [bn.from_mont(x, mod) for x in xs]
    assert _run(ctx, field, 6, xs, ys) == [bn.mont_mul(x, x, mod) for x in xs]


@pytest.mark.parametrize("field,mod", [(0, bn.R), (1, bn.Q)])
def test_mul_unreduced_first_operand(ctx, field, mod):
    """the first operand may be any 256-bit value (src/ultra_groth.cpp:54-55 feeds one)."""
    rng = random.Random(77 + field)
    xs = [(1 << 256) - 1, (1 << 256) - 2, mod, mod + 1, 2 * mod, 5 * mod + 3] + [rng.randrange(1 << 256) for _ in range(3000)]
    ys = [rng.randrange(mod) for _ in xs]
    assert _run(ctx, field, 0, xs, ys) == [bn.mont_mul(x, y, mod) for x, y in zip(xs, ys)]


@pytest.mark.parametrize("field,mod", [(0, bn.R), (1, bn.Q)])
def test_fp64_assisted_product_bit_exact(ctx, field, mod):
    """fp_mul_dfma (csrc/ff_variants.cuh): x*y on the integer multiplier, the m*p half of the Montgomery reduction on the FP64 pipe (exact
    52-bit limb products from round-towards-zero FMAs).  Same values as Fr_rawMMul for every input, including the rare case where
    the dropped low columns borrow through bit 256 (inputs built so that x*y mod 2^256 = 2^256 - L with L small)."""
    rng = random.Random(4321 + field)
    edge = [0, 1, 2, mod - 1, mod - 2, (1 << 256) % mod, (1 << 255) % mod, mod >> 1, 0xffffffff, 1 << 32, (1 << 64) - 1, (1 << 224) - 1]
    xs, ys = [], []
    for e in edge:
        for e2 in edge:
            xs.append(e)
            ys.append(e2)
    for _ in range(60000):
        xs.append(rng.randrange(mod))
        ys.append(rng.randrange(mod))
    for _ in range(20000):  # any 256-bit first operand
        xs.append(rng.randrange(1 << 256))
        ys.append(rng.randrange(mod))
    for _ in range(20000):  # borrow-repair path: x = (2^256 - L) * y^-1 mod 2^256, L of every magnitude up to 2^222
        y = rng.randrange(mod) | 1
        L = rng.randrange(1, 1 << rng.randrange(1, 223))
        xs.append(((1 << 256) - L) * pow(y, -1, 1 << 256) % (1 << 256))
        ys.append(y)
    assert _run(ctx, field, 7, xs, ys) == [bn.mont_mul(x, y, mod) for x, y in zip(xs, ys)]


def test_empty_vector(ctx):
    a = ctx.upload(b"")
    ctx.field_vec_op(0, 0, a, a, a, 0)
    ctx.sync()
