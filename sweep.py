#!/usr/bin/env python
"""Standalone sweep (BASELINE.json configs[4]): G1 (or G2) MSM and Fr NTT at 2^k sizes, at 1 / 2 / 4 / 8 GPUs, beside the CPU.

    python sweep.py --msm 16,18,20,22,24,26 --ntt 16,18,20,22,24,26 [--g2] [--dist uniform|witness] [--c WINDOW_BITS] [--reps 5] [--cpu]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 sweep.py ...      (N GPUs)

Everything goes through the C ABI (include/ultragroth_b200.h) and is timed with CUDA events on the library's own stream.
  * MSM at N GPUs: sharded by POINT RANGE: rank r keeps bases and scalars [r*n/N, (r+1)*n/N) resident, computes its
    XYZZ partial sum, the N partials (128 B each) are all-gathered and summed on the host.  Time = max over ranks of
    (device time of the partial MSM) + the exchange, i.e. the latency of ONE n-point MSM on N GPUs (strong scaling).
  * NTT at N GPUs: one independent size-n transform per GPU (the H pipeline's a / b / c transforms are independent,
    src/groth16.cpp:110-140): aggregate elements/s (weak scaling).
  * --cpu: the oracle's C++/OpenMP MSM and NTT (oracle/cpu_prover.cpp, "port": generic limbs, not ffiasm assembly) on the same inputs,
    all host cores, rank 0 only and only at N = 1 -- and the GPU result must EQUAL the CPU result (affine bytes / sha256 of all
    outputs): a full-size bit-exact check at every size of the sweep.
Prints one JSON line per (op, size).  Bases are k_i*G for seeded random k_i (GPU fixed-base generator), scalars seeded numpy bytes
below 2^253 ("uniform") or the witness-like mix of SURVEY.md 8(d).
"""
import argparse
import hashlib
import json
import os
import sys
import time

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
    if n:
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


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--msm", default="16,18,20,22")
    ap.add_argument("--ntt", default="16,20,22,24")
    ap.add_argument("--g2", action="store_true")
    ap.add_argument("--dist", default="uniform")
    ap.add_argument("--c", type=int, default=0)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--cpu", action="store_true", help="time the CPU oracle beside the GPU (N = 1 only) and require equal results")
    ap.add_argument("--cpu-max", type=int, default=24, help="largest log2 size the CPU column is run at")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import ultragroth_b200 as ug
    from ultragroth_b200 import sharded
    from ultragroth_b200._lib import lib
    L = lib()
    ctx = ug.Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local))
    group = 2 if args.g2 else 1
    cpu = args.cpu and world == 1
    threads = host_threads()

    def say(obj):
        if rank == 0:
            print(json.dumps(obj), flush=True)

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    for k in [int(x) for x in args.msm.split(",") if x]:
        n = 1 << k
        lo, hi = sharded.shard_range(n, world, rank)
        cnt = hi - lo
        bases_h = make_bases(ctx, L, group, cnt, seed=(0xB200 + k) * 64 + rank)
        bases = ctx.upload(bases_h)
        s1 = rand_scalars(cnt, seed=k * 64 + rank, dist=args.dist)
        d1 = ctx.upload(s1)

        def one_msm():
            part = ctx.msm(group, bases, d1, cnt, window_bits=args.c, xyzz=True)
            parts = sharded._allgather_bytes(part)
            return sharded.reduce_partials(b"".join(parts), group)

        r1 = one_msm()  # warm-up + result
        one_msm()
        # device time of the partial MSM (max over ranks) ...
        med, best = timed(torch, stream, lambda: ctx.msm(group, bases, d1, cnt, window_bits=args.c, xyzz=True), args.reps)
        med = max_over_ranks(med)
        # ... and wall time of the whole sharded MSM including the exchange and the host reduction
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            one_msm()
        wall = max_over_ranks((time.perf_counter() - t0) / args.reps * 1e3)
        rec = {"op": f"msm_g{group}", "log2n": k, "n_gpus": world, "dist": args.dist, "c": args.c or "auto", "ms_device": round(med, 4),
               "ms_end_to_end": round(wall, 4), "points_per_s": n / (wall * 1e-3), "points_per_s_device": n / (med * 1e-3)}
        if cpu and k <= args.cpu_max:
            from oracle import cpu_prover as cp
            t0 = time.perf_counter()
            want = cp.msm(group, bases_h, s1, n, threads=threads)
            t_cpu = time.perf_counter() - t0
            rec["cpu"] = {"ms": round(t_cpu * 1e3, 2), "points_per_s": n / t_cpu, "threads": threads, "kind": "port (oracle/cpu_prover.cpp)",
                          "gpu_speedup": t_cpu * 1e3 / wall}
            rec["equals_cpu_result"] = bool(want == r1)
            if not rec["equals_cpu_result"]:
                raise SystemExit(f"sweep: the GPU MSM at 2^{k} differs from the CPU oracle")
        say(rec)
        bases.free()
        d1.free()
        del bases_h, s1

    for k in [int(x) for x in args.ntt.split(",") if x]:
        n = 1 << k
        v = rand_scalars(n, seed=77 + k + 1000 * rank)
        d = ctx.upload(v)
        ctx.ntt(d, k)
        got = hashlib.sha256(ctx.download(d)).hexdigest() if (cpu and k <= args.cpu_max) else None
        med, best = timed(torch, stream, lambda: ctx.ntt(d, k), args.reps)
        med = max_over_ranks(med)
        d3 = [ctx.upload(v) for _ in range(3)]
        ctx.coset_transform(d3, k)
        medc, bestc = timed(torch, stream, lambda: ctx.coset_transform(d3, k), args.reps)
        medc = max_over_ranks(medc)
        rec = {"op": "ntt_fr", "log2n": k, "n_gpus": world, "ms_median": round(med, 4), "elements_per_s": world * n / (med * 1e-3),
               "coset_transform_x3_ms": round(medc, 4), "algorithmic_GBps_per_gpu": 64.0 * n * (2 if k > 12 else 1) / (med * 1e-3) / 1e9,
               "what": "one independent transform per GPU (weak scaling)" if world > 1 else "one transform"}
        if cpu and k <= args.cpu_max:
            from oracle import cpu_prover as cp
            t0 = time.perf_counter()
            want = cp.ntt(v, k, threads=threads)
            t_cpu = time.perf_counter() - t0
            rec["cpu"] = {"ms": round(t_cpu * 1e3, 2), "elements_per_s": n / t_cpu, "threads": threads, "kind": "port (oracle/cpu_prover.cpp)",
                          "note": "includes the root table build, as one FFT<Fr> call on a fresh object would", "gpu_speedup": t_cpu * 1e3 / med}
            rec["equals_cpu_result"] = bool(hashlib.sha256(want).hexdigest() == got)
            if not rec["equals_cpu_result"]:
                raise SystemExit(f"sweep: the GPU NTT at 2^{k} differs from the CPU oracle")
        say(rec)
        d.free()
        for x in d3:
            x.free()
    if world == 1:
        for kind, name in ((0, "imad"), (1, "imad_wide"), (3, "imad_wide_carry_chain"), (4, "imad_wide_carry_out"), (2, "fq_modmul")):
            say({"op": "int_pipe", "kind": name, "ops_per_s": ctx.bench_int_pipe(kind)})
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
