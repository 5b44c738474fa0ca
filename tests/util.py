"""Shared helpers for the tests: byte encodings between oracle integers and the zkey/device layouts."""
import random

import numpy as np

from oracle import bn254 as bn

M = bn.MONT_R


def fe_bytes(x):
    return int(x).to_bytes(32, "little")


def fe_list_bytes(xs):
    return b"".join(int(x).to_bytes(32, "little") for x in xs)


def bytes_fe_list(b):
    return [int.from_bytes(b[i:i + 32], "little") for i in range(0, len(b), 32)]


def g1_bytes(P):
    if P is None:
        return b"\0" * 64
    return fe_bytes(P[0] * M % bn.Q) + fe_bytes(P[1] * M % bn.Q)


def g2_bytes(P):
    if P is None:
        return b"\0" * 128
    return b"".join(fe_bytes(v * M % bn.Q) for v in (P[0][0], P[0][1], P[1][0], P[1][1]))


def g1_from_bytes(b):
    x, y = (int.from_bytes(b[i:i + 32], "little") for i in (0, 32))
    if x == 0 and y == 0:
        return None
    return (x * bn.FQ_RINV % bn.Q, y * bn.FQ_RINV % bn.Q)


def g2_from_bytes(b):
    v = [int.from_bytes(b[i:i + 32], "little") for i in (0, 32, 64, 96)]
    if not any(v):
        return None
    v = [x * bn.FQ_RINV % bn.Q for x in v]
    return ((v[0], v[1]), (v[2], v[3]))


def chain_points(curve, gen, n, seed):
    """n distinct affine points start + i*step (cheap to generate: one affine add each)."""
    rng = random.Random(seed)
    start = curve.mul(gen, rng.randrange(1, bn.R))
    step = curve.mul(gen, rng.randrange(1, bn.R))
    # batch: walk in Jacobian, convert with one shared inversion per chunk
    F = curve.F
    out = []
    cur = start
    jac = []
    for _ in range(n):
        jac.append(cur)
        cur = curve.add(cur, step)
    # Montgomery batch inversion of Z
    zs = [P[2] for P in jac]
    pref = []
    acc = F.one
    for z in zs:
        pref.append(acc)
        acc = F.mul(acc, z)
    inv = F.inv(acc)
    for i in range(n - 1, -1, -1):
        zi = F.mul(inv, pref[i])
        inv = F.mul(inv, zs[i])
        zi2 = F.sqr(zi)
        X, Y, _ = jac[i]
        out.append((F.mul(X, zi2), F.mul(Y, F.mul(zi2, zi))))
    out.reverse()
    return out


def scalars_uniform(n, seed):
    rng = random.Random(seed)
    return [rng.randrange(bn.R) for _ in range(n)]


def scalars_witness_like(n, seed):
    """50% zero, 20% one, 10% < 2^32, 20% uniform."""
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        u = rng.random()
        if u < 0.5:
            out.append(0)
        elif u < 0.7:
            out.append(1)
        elif u < 0.8:
            out.append(rng.randrange(1 << 32))
        else:
            out.append(rng.randrange(bn.R))
    return out
