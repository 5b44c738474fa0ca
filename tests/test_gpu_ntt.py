"""GPU parity: Fr NTT through the C ABI vs the oracle's restatement of FFT<Fr> (bit-exact)."""
import random

import pytest

from oracle import bn254 as bn
from tests.util import bytes_fe_list, fe_list_bytes

pytestmark = pytest.mark.gpu
R = bn.R
M = bn.MONT_R


def _mont(xs):
    return [x * M % R for x in xs]


def _unmont(xs):
    return [x * bn.FR_RINV % R for x in xs]


@pytest.mark.parametrize("logn", [0, 1, 2, 3, 5, 8, 10, 12, 13, 14])
def test_ntt_forward_inverse_vs_oracle(ctx, logn):
    n = 1 << logn
    rng = random.Random(400 + logn)
    xs = [rng.randrange(R) for _ in range(n)]
    d = ctx.upload(fe_list_bytes(_mont(xs)))
    ctx.ntt(d, logn, inverse=False)
    got = _unmont(bytes_fe_list(ctx.download(d)))
    assert got == bn.ntt(xs)
    ctx.ntt(d, logn, inverse=True)
    assert _unmont(bytes_fe_list(ctx.download(d))) == xs
    d2 = ctx.upload(fe_list_bytes(_mont(xs)))
    ctx.ntt(d2, logn, inverse=True)
    assert _unmont(bytes_fe_list(ctx.download(d2))) == bn.ntt(xs, inverse=True)


@pytest.mark.parametrize("logn", [4, 10, 13])
def test_coset_transform_vs_oracle(ctx, logn):
    """ifft, multiply by omega_2n^i, fft (src/groth16.cpp:110-140) on three vectors in one launch."""
    n = 1 << logn
    rng = random.Random(500 + logn)
    vecs = [[rng.randrange(R) for _ in range(n)] for _ in range(3)]
    ds = [ctx.upload(fe_list_bytes(_mont(v))) for v in vecs]
    ctx.coset_transform(ds, logn)
    w2n = bn.fr_root(logn + 1)
    for v, d in zip(vecs, ds):
        co = bn.ntt(v, inverse=True)
        co = [x * pow(w2n, i, R) % R for i, x in enumerate(co)]
        assert _unmont(bytes_fe_list(ctx.download(d))) == bn.ntt(co)


@pytest.mark.parametrize("logn", [16, 20, 22])
def test_ntt_large_properties(ctx, logn):
    """Size-independent checks at sizes the Python oracle cannot transform in seconds:
    (1) NTT of a delta at position k is the geometric sequence omega^(k*i);
    (2) inverse(forward(x)) == x on random data; (3) linearity spot check via X[0] = sum(x)."""
    import numpy as np
    n = 1 << logn
    k = 12345 % n
    w = bn.fr_root(logn)
    delta = np.zeros((n, 32), dtype=np.uint8)
    delta[k] = np.frombuffer((M % R).to_bytes(32, "little"), dtype=np.uint8)
    d = ctx.upload(delta)
    ctx.ntt(d, logn)
    out = ctx.download(d)
    wk = pow(w, k, R)
    for i in [0, 1, 2, 3, n // 2, n - 1, 777 % n, (n // 3) | 1]:
        got = int.from_bytes(out[32 * i:32 * i + 32], "little") * bn.FR_RINV % R
        assert got == pow(wk, i, R), i
    rng = np.random.default_rng(logn)
    raw = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    raw[:, 31] &= 0x1f  # < 2^253 < r
    d = ctx.upload(raw)
    ctx.ntt(d, logn)
    fwd = ctx.download(d)
    x0 = int.from_bytes(fwd[:32], "little")
    ints = raw.reshape(n, 4, 8).view("<u8").reshape(n, 4).astype(object)
    total = sum(int(ints[:, j].sum()) << (64 * j) for j in range(4)) % R
    assert x0 == total
    ctx.ntt(d, logn, inverse=True)
    assert ctx.download(d) == raw.tobytes()


SCHEDULES = [
    {"UGB_NTT_LOGE": "12", "UGB_NTT_MAX_LOGC": "11", "UGB_NTT_THREADS": "512"},   # one 128 KiB tile per SM (the first design)
    {"UGB_NTT_LOGE": "6", "UGB_NTT_MIN_LOGC": "0", "UGB_NTT_THREADS": "32"},      # 64-element tiles: 3 passes at 2^13
    {"UGB_NTT_LOGE": "7", "UGB_NTT_MIN_LOGC": "3", "UGB_NTT_MAX_LOGC": "4"},      # forces an extra pass for wide columns
    {"UGB_NTT_LOGE": "11", "UGB_NTT_MIN_LOGC": "0", "UGB_NTT_MAX_LOGC": "0"},     # single-column strided tiles
]


@pytest.mark.parametrize("sched", SCHEDULES, ids=lambda s: "-".join(f"{k[8:]}{v}" for k, v in s.items()))
def test_every_pass_schedule_is_bit_exact(ctx, sched, monkeypatch):
    """The pass planner's knobs (tile size, column width, thread count; read per call) only change the schedule:
    forward, inverse and the fused coset transform stay bit-exact against the oracle at a multi-pass size."""
    for k, v in sched.items():
        monkeypatch.setenv(k, v)
    logn = 13
    n = 1 << logn
    rng = random.Random(900)
    vecs = [[rng.randrange(R) for _ in range(n)] for _ in range(3)]
    d = ctx.upload(fe_list_bytes(_mont(vecs[0])))
    ctx.ntt(d, logn)
    assert _unmont(bytes_fe_list(ctx.download(d))) == bn.ntt(vecs[0])
    ctx.ntt(d, logn, inverse=True)
    assert _unmont(bytes_fe_list(ctx.download(d))) == vecs[0]
    ds = [ctx.upload(fe_list_bytes(_mont(v))) for v in vecs]
    ctx.coset_transform(ds, logn)
    w2n = bn.fr_root(logn + 1)
    for v, dv in zip(vecs, ds):
        co = bn.ntt(v, inverse=True)
        co = [x * pow(w2n, i, R) % R for i, x in enumerate(co)]
        assert _unmont(bytes_fe_list(ctx.download(dv))) == bn.ntt(co)
