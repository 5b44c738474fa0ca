#!/usr/bin/env python
"""Per-shard device time of the three compute phases of the sharded prover, with all `world` shards living on ONE GPU
(sharded.prove_local): shows whether the shards carry equal work, without needing `world` GPUs.

    python tools/shard_phase_times.py [--world 8] [--logn 22] [--workload ultragroth|groth16]

Prints one line per shard: witness_msms / h_transform(all three vectors) / h_msm in ms (CUDA events on the shard's stream, median of
--reps), after one whole prove_local proof that is verified.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--logn", type=int, default=22)
    ap.add_argument("--workload", default="ultragroth")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--kernels", action="store_true", help="per-kernel device time of shard 0's witness_msms phase (profile mode: CUDA events per launch)")
    args = ap.parse_args()
    import torch
    import ultragroth_b200 as ug
    from ultragroth_b200 import sharded
    from ultragroth_b200 import verifier as uv
    from bench import lookup_params
    ultra = args.workload == "ultragroth"
    ctx = ug.Context(0)
    lookup_bits, chunks = lookup_params(args.logn)
    zk, wt, vkj = ctx.synth(args.logn, seed=0xB200 + args.logn, ultra=ultra, lookup_bits=lookup_bits, n_chunks=chunks)
    shards = [sharded.ShardedProver(zk, args.workload, rank=r, world=args.world) for r in range(args.world)]
    rk = (1 << 200) + 0xB200
    res = sharded.prove_local(shards, wt, rk if ultra else None)
    ok = uv.verify(args.workload, res[0][0], res[0][1], vkj)
    print("prove_local over %d shards on one GPU: verified = %s" % (args.world, bool(ok)), flush=True)
    if not ok:
        raise SystemExit(1)
    for r, sp in enumerate(shards):
        lay = sp.layout()
        st = torch.cuda.ExternalStream(lay.stream, device=torch.device("cuda", lay.device))
        out = {}
        for name, fn in (("witness_msms", sp.witness_msms), ("h_transform(7)", lambda: sp.h_transform(7)), ("h_msm", sp.h_msm)):
            ts = []
            with torch.cuda.stream(st):
                for _ in range(args.reps):
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    fn()
                    b.record()
                    b.synchronize()
                    ts.append(a.elapsed_time(b))
            out[name] = sorted(ts)[len(ts) // 2]
        sp.finish()
        print("shard %d/%d: " % (r, args.world) + ", ".join("%s %.2f ms" % kv for kv in out.items()), flush=True)
    if args.kernels:
        import ctypes
        from ultragroth_b200._lib import lib
        L = lib()
        sp = shards[0]
        sp.witness_msms()
        sp.finish()
        L.ugb_prover_set_profile(sp.handle, 1)
        reps = 3
        for _ in range(reps):
            sp.witness_msms()
            sp.finish()
        n = L.ugb_prover_kernel_stats(sp.handle, None, 0)
        buf = ctypes.create_string_buffer(int(n) + 16)
        L.ugb_prover_kernel_stats(sp.handle, buf, len(buf))
        L.ugb_prover_set_profile(sp.handle, 0)
        rows = []
        for line in buf.value.decode().splitlines():
            name, cnt, ms = line.split("\t")
            rows.append((float(ms) / reps, int(cnt) // reps, name))
        print("shard 0 witness_msms by kernel (ms per phase, launches): total %.2f ms" % sum(r[0] for r in rows))
        for ms, cnt, name in sorted(rows, reverse=True):
            print("  %-28s %3d  %.3f" % (name, cnt, ms))
    for sp in shards:
        sp.close()


if __name__ == "__main__":
    main()
