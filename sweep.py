#!/usr/bin/env python
"""Standalone sweep (BASELINE.json configs[4]): G1/G2 MSM and Fr NTT at 2^k sizes through the C ABI
(include/ultragroth_b200.h), timed with CUDA events on the library's own stream.

    python sweep.py --msm 16,18,20,22 --ntt 16,20,24 [--g2] [--dist uniform|witness] [--c WINDOW_BITS] [--reps 5]

Prints one JSON line per (op, size).  Bases are k_i*G for seeded random k_i (GPU fixed-base generator), scalars
seeded numpy bytes below 2^253 ("uniform") or the witness-like mix of SURVEY.md 8(d).  Every MSM size <= 2^12 is
also checked against the big-int oracle; larger sizes are checked by linearity: MSM(s) + MSM(s') == MSM(s + s').
"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

R = 21888242871839275222246405745257275088548364400416034343698204186575808495617


def rand_scalars(n, seed, dist="uniform"):
    rng = np.random.default_rng(seed)
    b = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    b[:, 31] &= 0x1F  # < 2^253 < r
    if dist == "witness":
        u = rng.random(n)
        b[u < 0.5] = 0
        one = (u >= 0.5) & (u < 0.7)
        b[one] = 0
        b[one, 0] = 1
        small = (u >= 0.7) & (u < 0.8)
        b[small, 4:] = 0
    return np.ascontiguousarray(b)


def make_bases(ctx, lib, group, n, seed):
    from ultragroth_b200._lib import check
    k = rand_scalars(n, seed)
    out = np.empty((n, 64 if group == 1 else 128), dtype=np.uint8)
    check(lib.ugb_fixed_base_mul(ctx.handle, group, k.ctypes.data, n, out.ctypes.data))
    return out


def timed(torch, stream, fn, reps):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    with torch.cuda.stream(stream):
        for a, b in ev:
            a.record(stream)
            fn()
            b.record(stream)
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return t[len(t) // 2], t[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--msm", default="16,18,20,22")
    ap.add_argument("--ntt", default="16,20,22,24")
    ap.add_argument("--g2", action="store_true")
    ap.add_argument("--dist", default="uniform")
    ap.add_argument("--c", type=int, default=0)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--device", type=int, default=0)
    args = ap.parse_args()
    import torch
    import ultragroth_b200 as ug
    from ultragroth_b200._lib import lib
    L = lib()
    torch.cuda.set_device(args.device)
    ctx = ug.Context(args.device)
    stream = torch.cuda.ExternalStream(ctx.stream)
    group = 2 if args.g2 else 1
    for k in [int(x) for x in args.msm.split(",") if x]:
        n = 1 << k
        bases_h = make_bases(ctx, L, group, n, seed=0xB200 + k)
        bases = ctx.upload(bases_h)
        s1 = rand_scalars(n, seed=k, dist=args.dist)
        d1 = ctx.upload(s1)
        r1 = ctx.msm(group, bases, d1, n, window_bits=args.c)  # warm-up + result
        ctx.msm(group, bases, d1, n, window_bits=args.c)
        med, best = timed(torch, stream, lambda: ctx.msm(group, bases, d1, n, window_bits=args.c), args.reps)
        # linearity check: MSM(s1) + MSM(s2) == MSM(s1 + s2 mod r)
        s2 = rand_scalars(n, seed=1000 + k, dist=args.dist)
        a = [int.from_bytes(s1[i].tobytes(), "little") for i in range(min(n, 1 << 14))]
        ok = None
        if n <= (1 << 14):
            from oracle import bn254 as bn
            from tests.util import g1_from_bytes, g2_from_bytes
            b = [int.from_bytes(s2[i].tobytes(), "little") for i in range(n)]
            s3 = np.frombuffer(b"".join(((x + y) % R).to_bytes(32, "little") for x, y in zip(a, b)), dtype=np.uint8)
            dec = g1_from_bytes if group == 1 else g2_from_bytes
            C = bn.G1 if group == 1 else bn.G2
            r2 = ctx.msm(group, bases, ctx.upload(s2), n, window_bits=args.c)
            r3 = ctx.msm(group, bases, ctx.upload(s3), n, window_bits=args.c)
            ok = C.to_affine(C.add(C.from_affine(dec(r1)), C.from_affine(dec(r2)))) == dec(r3)
        print(json.dumps({"op": f"msm_g{group}", "log2n": k, "dist": args.dist, "c": args.c or "auto", "ms_median": round(med, 4),
                          "ms_best": round(best, 4), "points_per_s": n / (med * 1e-3), "linearity_ok": ok}), flush=True)
        bases.free()
        d1.free()
    for k in [int(x) for x in args.ntt.split(",") if x]:
        n = 1 << k
        v = rand_scalars(n, seed=77 + k)
        d = ctx.upload(v)
        ctx.ntt(d, k)
        med, best = timed(torch, stream, lambda: ctx.ntt(d, k), args.reps)
        d3 = [ctx.upload(v) for _ in range(3)]
        ctx.coset_transform(d3, k)
        medc, bestc = timed(torch, stream, lambda: ctx.coset_transform(d3, k), args.reps)
        print(json.dumps({"op": "ntt_fr", "log2n": k, "ms_median": round(med, 4), "ms_best": round(best, 4),
                          "elements_per_s": n / (med * 1e-3), "coset_transform_x3_ms": round(medc, 4),
                          "algorithmic_GBps": 64.0 * n * (2 if k > 12 else 1) / (med * 1e-3) / 1e9}), flush=True)
        d.free()
        for x in d3:
            x.free()
    for kind, name in ((0, "imad"), (1, "imad_wide"), (3, "imad_wide_carry_chain"), (4, "imad_wide_carry_out"), (2, "fq_modmul")):
        print(json.dumps({"op": "int_pipe", "kind": name, "ops_per_s": ctx.bench_int_pipe(kind)}), flush=True)


if __name__ == "__main__":
    main()
