#!/usr/bin/env python
"""Headline benchmark: proofs/s (and prove latency) on a synthetic 2^22-constraint BN254 circuit.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload ultragroth|groth16] [--logn 22] [--impl reference]

One process per GPU (torchrun for N > 1).  Every rank holds a resident prover (zkey bases, coefficients and
twiddles in HBM) and proves independent witnesses: proofs shard across GPUs with no data-path collective
(weak scaling).  `value` is the device-resident leg (witness already in HBM), `e2e` goes through the
reference's C API (`*_prover_prove`, include/prover.h) with host buffers.  The CPU baseline is the oracle's
C++/OpenMP restatement of the reference prover (oracle/cpu_prover.cpp) on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("UGB_BENCH_WORKLOAD", "ultragroth"), choices=["ultragroth", "groth16"])
    ap.add_argument("--logn", type=int, default=int(os.environ.get("UGB_BENCH_LOGN", "22")))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-shard", action="store_true", help="skip the sharded single-proof leg that --gpus N > 1 adds to the line")
    ap.add_argument("--shard", action="store_true",
                    help="one proof at a time over ALL ranks: MSMs sharded by point range (SURVEY 8(e)) instead of replicas")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """nvidia-smi SM clocks and throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = threading.Event()

    def run(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def workload_name(args):
    if args.workload == "ultragroth":
        return f"synthetic UltraGroth circuit (chain + 2^16-entry range lookup, 2^{args.logn - 2} chunks), 2^{args.logn} constraints"
    return f"synthetic Groth16 chain circuit, 2^{args.logn} constraints"


def lookup_params(logn):
    lookup_bits = min(16, max(2, logn - 6))
    chunks = 1 << max(2, logn - 2)
    return lookup_bits, chunks


def host_threads():
    """Cores this process may run on.  torchrun exports OMP_NUM_THREADS=1, which must not turn the CPU arm into a 1-thread run:
    the oracle takes its thread count explicitly (num_threads clauses), never from the OpenMP environment."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_inputs(args):
    """The benchmark's zkey / witness made on the CPU (oracle/cpu_synth.cpp: same seed => the same bytes as the GPU generator,
    tests/test_gpu_synth.py), cached under /tmp for the other runs on this box.  Does not load the product library."""
    import numpy as np
    from oracle import cpu_prover
    ultra = args.workload == "ultragroth"
    lookup_bits, chunks = lookup_params(args.logn)
    key = f"ugb_synth_{args.workload}_{args.logn}_{lookup_bits}_{chunks}"
    zp, wp = f"/tmp/{key}.zkey.npy", f"/tmp/{key}.wtns.npy"
    t0 = time.time()
    try:
        if os.path.exists(zp) and os.path.exists(wp):
            return np.load(zp), np.load(wp), {"source": "cache", "seconds": round(time.time() - t0, 2)}
    except Exception:
        pass
    zk, wt, _ = cpu_prover.synth(args.logn, seed=0xB200 + args.logn, ultra=ultra, lookup_bits=lookup_bits, n_chunks=chunks, threads=host_threads())
    gen_s = time.time() - t0
    try:
        tmp = f"{zp}.{os.getpid()}.npy"
        np.save(tmp, zk); os.replace(tmp, zp)
        tmp = f"{wp}.{os.getpid()}.npy"
        np.save(tmp, wt); os.replace(tmp, wp)
    except Exception:
        pass
    return zk, wt, {"source": "oracle/cpu_synth.cpp", "seconds": round(gen_s, 2)}


def cpu_prove_timed(args, zk, wt, steps, budget_s):
    """Times REAL proofs of the quoted workload on the host cores with the oracle's C++/OpenMP restatement of the reference prover
    (oracle/cpu_prover.cpp): as many of `steps` as fit `budget_s` (at least one).  -> (result dict, seconds per proof list)"""
    from oracle import cpu_prover
    threads = host_threads()
    times, stage = [], None
    t_begin = time.perf_counter()
    for i in range(max(1, steps)):
        t0 = time.perf_counter()
        res = cpu_prover.prove(zk, wt, 12345 + i, 67890 + i, 424242 + i, threads=threads)
        times.append(time.perf_counter() - t0)
        stage = res["times_ms"]
        elapsed = time.perf_counter() - t_begin
        if elapsed + times[-1] > budget_s:
            break
    per = sum(times) / len(times)
    out = {"value": 1.0 / per, "unit": "proofs/s", "cores": threads, "kind": "port", "steps_timed": len(times),
           "seconds_per_proof": [round(t, 3) for t in times],
           "sample": f"{len(times)} whole {args.workload} proof(s) of the quoted workload (2^{args.logn} constraints, same seed and files as the "
                     f"GPU arm), wall clock around cpu_prove on {threads} threads; nothing extrapolated",
           "stage_ms": {"msm": stage[0], "spmv": stage[1], "fft": stage[2], "prove_without_root_tables": stage[3], "root_tables": stage[4]},
           "note": "C++17/OpenMP restatement of src/groth16.cpp / src/ultra_groth.cpp on generic 64-bit limbs (oracle/cpu_prover.cpp), "
                   "not ffiasm's x86-64 assembly: the reference's own engine is absent from the reference tree (empty submodule, no nasm) "
                   "and cannot be built here, so real rapidsnark would be faster than this arm"}
    return out


def run_reference(args):
    """--impl reference: the CPU prover on this box's host cores, same config / metric / unit as the GPU arm.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.time()
    zk, wt, how = cpu_inputs(args)
    # each step is one whole proof; the run is bounded to ~2 minutes of proving (the driver asks for K steps, a 2^22 proof takes
    # 10-20 s on a typical host), and the number of steps really timed is reported
    budget = float(os.environ.get("UGB_BENCH_CPU_BUDGET_S", "100"))
    t1 = time.time()
    res = cpu_prove_timed(args, zk, wt, args.steps, budget)
    t_timed = time.time() - t1
    per_ms = 1000.0 / res["value"]
    line = {
        "impl": "reference", "metric": "proofs_per_sec", "value": res["value"], "unit": "proofs/s", "n_gpus": args.gpus,
        "steps": res["steps_timed"], "steps_requested": args.steps, "steps_timed": res["steps_timed"], "warmup": 0,
        "ms_per_step": per_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u256 (4x64-bit Montgomery limbs)", "data": "synthetic", "config": {"workload": workload_name(args)},
        "cpu_baseline": res,
        "e2e": {"value": res["value"], "unit": "proofs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "inputs": how, "timed_s": round(t_timed, 2), "wall_s": round(time.time() - t0, 2),
        "note": "CPU arm: one process, all host cores, whatever --gpus says (N GPUs are compared with the same one host)",
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ["UGB_DEVICE"] = str(local)
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    import ultragroth_b200 as ug
    from ultragroth_b200 import prover as up
    from oracle import groth16 as og

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.shard:
        return main_sharded(args, torch, dist, world, rank, local)
    ctx = ug.Context(local)
    ultra = args.workload == "ultragroth"
    lookup_bits, chunks = lookup_params(args.logn)
    t_setup = time.time()
    zkey, wtns, vkj = ctx.synth(args.logn, seed=0xB200 + args.logn, ultra=ultra, lookup_bits=lookup_bits, n_chunks=chunks)
    t_setup = time.time() - t_setup
    # IMAD.WIDE issues at half the IMAD rate on sm_100 (32 lanes/clk/SM); three differently-shaped chains are
    # measured (plain, carry-linked .X rows as in fp_mul, carry-out only) and the best is taken as the pipe's peak
    wide_kinds = {"plain": ctx.bench_int_pipe(1), "carry_chain": ctx.bench_int_pipe(3), "carry_out": ctx.bench_int_pipe(4)}
    imad_wide = max(wide_kinds.values())
    imad = ctx.bench_int_pipe(0)
    modmul = ctx.bench_int_pipe(2)
    t_create = time.time()
    prover = up.Prover(zkey, "ultragroth" if ultra else "groth16")
    t_create = time.time() - t_create
    zkey_bytes = int(zkey.nbytes)
    create_ms = prover.create_times()
    # kept on the host: the CPU baseline proves the very same files (N = 1), the sharded provers are created from them (N > 1)

    # pinned host witness: what the C API is handed every step
    wt = torch.from_numpy(wtns).pin_memory()
    wt_np = wt.numpy()

    # ---- warm-up (also the correctness gate: the proof must verify under the repo's verifier equation) ----
    proof_json = public_json = None
    for _ in range(max(args.warmup, 3)):
        proof_json, public_json = prover.prove(wt_np)
    verified = None
    if rank == 0 and not args.no_verify:
        # the product's own verifier (include/verifier.h, host C++) is the gate; the oracle's verifier cross-checks it
        from ultragroth_b200 import verifier as uv
        verified = uv.verify("ultragroth" if ultra else "groth16", proof_json, public_json, vkj)
        proof = og.proof_from_json(proof_json)
        public = [int(x) for x in json.loads(public_json)] if public_json != "null" else []
        if ultra:
            verified = verified and bool(og.ultragroth_verify(og.vk_from_ultragroth_json(vkj), proof, public))
        else:
            verified = verified and bool(og.groth16_verify(og.vk_from_snarkjs_json(vkj), proof, public))
        if not verified:
            raise SystemExit("bench: the proof does not verify -- refusing to report a number")

    sampler = ClockSampler(local)
    sampler.start()

    # ---- device-resident leg: K proofs on the witness already in HBM ----
    prover.set_profile(True)
    prover.set_timing(True)
    launches0 = prover.kernel_launches
    pstream = torch.cuda.ExternalStream(prover.stream, device=torch.device("cuda", local))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.profiler.start()  # ncu --profile-from-start off: only the timed proofs are captured
    t0 = time.perf_counter()
    ev0.record(pstream)
    res_proof = None
    for _ in range(args.steps):
        res_proof = prover.prove_resident()  # a WHOLE proof each (both rounds, challenge, lookup fill, MSMs, H, assembly, JSON)
    ev1.record(pstream)
    barrier()
    t_wall = time.perf_counter() - t0
    torch.cuda.profiler.stop()
    # CUDA events on the prover's own stream; every prove_resident call returns only when its proof is assembled on the host, so the
    # interval covers the host tail of each proof as well
    t_dev = ev0.elapsed_time(ev1) * 1e-3
    if rank == 0 and not args.no_verify:
        from ultragroth_b200 import verifier as uv
        if not uv.verify("ultragroth" if ultra else "groth16", res_proof[0], res_proof[1], vkj):
            raise SystemExit("bench: the device-resident proof does not verify -- refusing to report a number")
    launches = prover.kernel_launches - launches0
    stats = prover.kernel_stats()
    stage = prover.last_times()
    prover.set_profile(False)
    prover.set_timing(False)

    # ---- end-to-end leg: the reference-facing C API with host buffers (H2D witness, D2H results inside) ----
    barrier()
    t0 = time.perf_counter()
    ev0.record(pstream)
    for _ in range(args.steps):
        proof_json, public_json = prover.prove(wt_np)
    ev1.record(pstream)
    barrier()
    t_e2e_wall = time.perf_counter() - t0
    # the C API returns after the host has assembled the proof JSON: wall clock covers what the events cannot
    t_e2e_seq = max(ev0.elapsed_time(ev1) * 1e-3, t_e2e_wall)
    # ... and the batch entry point (ugb_prover_prove_batch): the same K proofs from the same pinned host buffers, but the upload of
    # witness k+1 runs under proof k and the host tail of proof k under proof k+1.  Every step still copies its witness host -> device
    # and its results device -> host inside the timed region.
    up.prove_batch(prover, [wt_np] * 2)  # untimed: the batch path's own buffers (second witness buffer, second result slots) get allocated
    barrier()
    t0 = time.perf_counter()
    ev0.record(pstream)
    batch_out = up.prove_batch(prover, [wt_np] * args.steps)
    ev1.record(pstream)
    barrier()
    t_e2e = max(ev0.elapsed_time(ev1) * 1e-3, time.perf_counter() - t0)
    if rank == 0 and not args.no_verify:
        from ultragroth_b200 import verifier as uv
        if not all(uv.verify("ultragroth" if ultra else "groth16", pj, uj, vkj) for pj, uj in (batch_out[0], batch_out[-1])):
            raise SystemExit("bench: a proof of the pipelined batch does not verify -- refusing to report a number")
    sampler.stop_flag.set()
    sampler.join(timeout=2)

    if world > 1:
        t = torch.tensor([t_dev, t_e2e, t_e2e_seq], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_dev, t_e2e, t_e2e_seq = float(t[0]), float(t[1]), float(t[2])

    # ---- the north star's multi-GPU split, measured in the same run: ONE proof at a time over all N ranks ----
    shard_rec = None
    if world > 1 and not args.no_shard:
        shard_rec = sharded_leg(args, torch, dist, world, rank, local, zkey, wt_np, vkj, ultra, barrier)

    if rank == 0:
        n = 1 << args.logn
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"
        # dominant kernel = largest share of event-timed device time
        kstats = {k: v for k, v in stats.items() if not k.startswith("#")}
        tot_ms = sum(ms for _, ms in kstats.values()) or 1.0
        top = sorted(kstats.items(), key=lambda kv: -kv[1][1])
        kern_share = {k: round(v[1] / tot_ms, 4) for k, v in top[:8]}
        acc_l, acc_ms = stats.get("k_accumulate<G1>", (0, 0.0))
        # G1 bucket accumulation: every sorted digit entry = one 4-byte index read + one 64-byte affine point gather,
        # and 10 Montgomery products (madd-2008-s) = 1360 IMAD.WIDE
        nvars = int.from_bytes(wtns[60:64].tobytes(), "little")
        line = {
            "metric": "proofs_per_sec", "value": args.steps * world / t_dev, "unit": "proofs/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": 1000.0 * t_dev / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u256 (8x32-bit Montgomery limbs, BN254 Fr/Fq)", "data": "synthetic",
            "config": {"workload": workload_name(args), "nVars": nvars, "domainSize": n, "parallelism": f"replica x{world}",
                       "l2": "inputs exceed L2 (zkey bases %.2f GB resident in HBM)" % (zkey_bytes / 1e9),
                       "blinding": "random r, s (and r_k)"},
            "e2e": {"value": args.steps * world / t_e2e, "unit": "proofs/s", "ms_per_proof": 1000.0 * t_e2e / args.steps,
                    "h2d_bytes_per_step": int(wtns.nbytes), "d2h_bytes_per_step": 6 * 255 * 128 + 255 * 256 + 32 * 2,
                    "api": "ugb_prover_prove_batch (K witnesses from pinned host buffers, pipelined)",
                    "sequential": {"value": args.steps * world / t_e2e_seq, "ms_per_proof": 1000.0 * t_e2e_seq / args.steps,
                                   "api": "ultra_groth_prover_prove" if ultra else "groth16_prover_prove"}},
            "gpu_launches": int(launches),
            "clocks": sampler.summary(),
            "verified": verified,
            "stage_ms": {k: round(v, 3) for k, v in stage.items()},
            "kernel_share": kern_share,
            "int_pipe_peak": {"imad_per_s": imad, "imad_wide_per_s": imad_wide, "imad_wide_kinds": wide_kinds, "fq_modmul_per_s": modmul},
            "wall_ms_per_step": 1000.0 * t_wall / args.steps,
            "setup_s": round(t_setup, 2), "create_s": round(t_create, 2), "create_ms": {k: round(v, 1) for k, v in create_ms.items()},
            "inf_points": prover.inf_points(),
        }
        # second half of BASELINE.json's metric: G1 MSM points/s, taken from the proof's own H MSM (domainSize dense full-width
        # scalars against the resident table, sort + accumulate + reduce) and the A MSM (nVars witness scalars); whole job = x world
        if stage.get("msm_h", 0) > 0 and stage.get("msm_a", 0) > 0:
            line["g1_msm"] = {
                "h_points_per_s": world * n / ((stage["msm_h"] + stage["sort_h"]) * 1e-3), "h_points": n, "h_ms": stage["msm_h"] + stage["sort_h"],
                "a_points_per_s": world * nvars / (stage["msm_a"] * 1e-3), "a_points": nvars, "a_ms": stage["msm_a"],
                "note": "per-proof MSMs over resident window tables (window picked by msm_pick_config: c=20 at 2^22); standalone ad-hoc-base sweep: profiles/r01_sweep_g1.jsonl"}
        line["prove_latency_ms"] = {"resident": 1000.0 * t_dev / args.steps, "e2e_single_call": 1000.0 * t_e2e_seq / args.steps,
                                    "e2e_pipelined_per_proof": 1000.0 * t_e2e / args.steps}
        line["roofline"], line["roofline_ntt"] = rooflines(stats, args, nvars, n, hbm_peak, peak_src, imad_wide, ultra, chunks, lookup_bits)
        if shard_rec is not None:
            line["sharded"] = shard_rec
        if not args.no_cpu and world == 1:
            try:
                line["cpu_baseline"] = cpu_prove_timed(args, zkey, wtns, 1, 0.0)
            except Exception as e:  # the baseline must never take the GPU number down with it
                line["cpu_baseline"] = {"error": str(e)[:200]}
        print(json.dumps(line), flush=True)
    prover.close()
    if world > 1:
        dist.destroy_process_group()


def sharded_leg(args, torch, dist, world, rank, local, zkey, wt_np, vkj, ultra, barrier):
    """Single-proof latency with everything sharded over the N ranks (ultragroth_b200/sharded.py prove2): every rank keeps 1/N of each
    base table, uploads 1/N of the witness (all-gathered over NVLink), the a / b / c transforms run on three ranks and each rank gets
    its range of a', b', c' for its share of the H MSM; the 896-byte partial sums are gathered and rank 0 assembles.  Also the
    metric's "G1 MSM points/s at N GPUs": the sharded H MSM (domainSize dense scalars: digit sort + accumulate + reduce per rank)."""
    from ultragroth_b200 import sharded
    from ultragroth_b200 import verifier as uv
    t0 = time.time()
    sp = sharded.ShardedProver(zkey, "ultragroth" if ultra else "groth16", rank=rank, world=world)
    t_create = time.time() - t0
    comm = sharded.DistComm(sp)
    rk = (1 << 200) + 0xB200  # common r_k for UltraGroth (any value < 2^248; fixed so that every rank agrees)
    res = None
    for _ in range(max(args.warmup, 3)):
        res = sp.prove2(wt_np, rk=rk if ultra else None, comm=comm)
    ok = None
    if rank == 0 and not args.no_verify:
        ok = bool(uv.verify("ultragroth" if ultra else "groth16", res[0], res[1], vkj))
        if not ok:
            raise SystemExit("bench: the sharded proof does not verify -- refusing to report a number")
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sp.prove2(wt_np, rk=rk if ultra else None, comm=comm)
    barrier()
    t = time.perf_counter() - t0
    # G1 MSM points/s: the H MSM phase alone, timed with CUDA events on the prover's stream (h already in HBM from the last proof)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    barrier()
    with comm.stream_scope():
        ev0.record()
        for _ in range(reps):
            sp.h_msm()
        ev1.record()
    ev1.synchronize()
    t_h = ev0.elapsed_time(ev1) * 1e-3 / reps
    tt = torch.tensor([t, t_h], dtype=torch.float64, device="cuda")
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t, t_h = float(tt[0]), float(tt[1])
    # where the latency goes: two TRACED proofs after the timed ones (the stream is drained at every seam, so the phases are
    # serialised and sum to more than ms_per_proof); the max over ranks per phase shows which phase the slowest shard holds up
    trace = {}
    for _ in range(2):
        sp.prove2(wt_np, rk=rk if ultra else None, comm=comm, trace=trace)
    names = ["upload_slice", "all_gather_witness", "round_msm_and_exchange", "lookup", "h_transform", "witness_msms", "scatter_wait",
             "h_msm_and_finish", "exchange_partials", "assemble"]
    tl = torch.tensor([trace.get(k, 0.0) / 2 for k in names], dtype=torch.float64, device="cuda")
    tl_max = tl.clone()
    dist.all_reduce(tl_max, op=dist.ReduceOp.MAX)
    n = 1 << args.logn
    rec = {"scaling": "strong", "ms_per_proof": 1000.0 * t / args.steps, "proofs_per_s": args.steps / t, "verified": ok,
           "parallelism": f"bases, witness upload, H pipeline and MSMs sharded x{world}; one proof at a time",
           "g1_msm_h": {"points": n, "ms": 1000.0 * t_h, "points_per_s": n / t_h,
                        "what": "H MSM of the proof sharded by point range: k_h_final + digit sort + bucket accumulation + reduction per rank, max over ranks"},
           "create_s": round(t_create, 2),
           "timeline_rank0_ms": {k: round(float(v), 3) for k, v in zip(names, tl.tolist()) if v > 0},
           "timeline_max_over_ranks_ms": {k: round(float(v), 3) for k, v in zip(names, tl_max.tolist()) if v > 0}}
    sp.close()
    return rec


def main_sharded(args, torch, dist, world, rank, local):
    """Single-proof latency with the MSMs sharded by point range over all ranks (strong scaling): every rank keeps 1/N of
    each base array resident, the H pipeline is replicated, one all-gather of the partial sums per phase, rank 0 assembles."""
    import secrets
    import ultragroth_b200 as ug
    from ultragroth_b200 import sharded
    from oracle import groth16 as og
    ctx = ug.Context(local)
    ultra = args.workload == "ultragroth"
    lookup_bits, chunks = lookup_params(args.logn)
    t_setup = time.time()
    zkey, wtns, vkj = ctx.synth(args.logn, seed=0xB200 + args.logn, ultra=ultra, lookup_bits=lookup_bits, n_chunks=chunks)
    t_setup = time.time() - t_setup
    t_create = time.time()
    sp = sharded.ShardedProver(zkey, "ultragroth" if ultra else "groth16", rank=rank, world=world)
    t_create = time.time() - t_create
    zkey_bytes = int(zkey.nbytes)
    del zkey
    wt = torch.from_numpy(wtns).pin_memory().numpy()
    rk = (1 << 200) + 0xB200  # common r_k for UltraGroth (any value < 2^248; fixed so that every rank agrees)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    comm = sharded.DistComm(sp)
    res = None
    for _ in range(max(args.warmup, 3)):
        res = sp.prove2(wt, rk=rk if ultra else None, comm=comm)
    verified = None
    if rank == 0 and not args.no_verify:
        proof = og.proof_from_json(res[0])
        public = [int(x) for x in json.loads(res[1])] if res[1] != "null" else []
        verified = bool(og.ultragroth_verify(og.vk_from_ultragroth_json(vkj), proof, public) if ultra
                        else og.groth16_verify(og.vk_from_snarkjs_json(vkj), proof, public))
        if not verified:
            raise SystemExit("bench: the sharded proof does not verify -- refusing to report a number")
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0 = time.perf_counter()
    ev0.record()
    for _ in range(args.steps):
        sp.prove2(wt, rk=rk if ultra else None, comm=comm)
    ev1.record()
    barrier()
    t = max(time.perf_counter() - t0, ev0.elapsed_time(ev1) * 1e-3)
    sampler.stop_flag.set()
    sampler.join(timeout=2)
    if world > 1:
        tt = torch.tensor([t], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t = float(tt[0])
    if rank == 0:
        nvars = int.from_bytes(wtns[60:64].tobytes(), "little")
        line = {"metric": "proofs_per_sec", "value": args.steps / t, "unit": "proofs/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": 1000.0 * t / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "u256 (8x32-bit Montgomery limbs, BN254 Fr/Fq)", "data": "synthetic",
                "config": {"workload": workload_name(args), "nVars": nvars, "domainSize": 1 << args.logn,
                           "parallelism": f"bases, witness upload, H pipeline and MSMs sharded x{world}",
                           "l2": "inputs exceed L2 (zkey bases %.2f GB over %d GPUs)" % (zkey_bytes / 1e9, world)},
                "e2e": {"value": args.steps / t, "unit": "proofs/s", "ms_per_proof": 1000.0 * t / args.steps,
                        "h2d_bytes_per_step": int(wtns.nbytes), "d2h_bytes_per_step": int(sp.partial_size),
                        "api": "ugb_*_prove_partial + all_gather + ugb_*_prove_assemble"},
                "gpu_launches": None, "clocks": sampler.summary(), "verified": verified,
                "setup_s": round(t_setup, 2), "create_s": round(t_create, 2)}
        print(json.dumps(line), flush=True)
    sp.close()
    if world > 1:
        dist.destroy_process_group()


def rooflines(stats, args, nvars, n, hbm_peak, peak_src, imad_wide_peak, ultra, chunks, lookup_bits):
    """Roofline entries for the dominant kernel (G1 bucket accumulation) and for the NTT passes.

    k_accumulate<G1> is bound by the 32-bit integer multiplier (MSM roofline = 136 * modmuls / time / imad_peak):
    per sorted digit entry one mixed addition = 10 Montgomery products = 1360 IMAD.WIDE, against 4 B (entry) + 64 B (affine point
    gather) of algorithmic HBM traffic.  The top-level object is therefore the integer-pipe form; the HBM form the bench contract
    asks for rides along as `hbm`.  The entry count is the library's own counter over the timed launches; the peak is measured
    in this very run (ugb_bench_int_pipe) and the arithmetic peak (SMs x 32 lanes x clock) is printed beside it.
    k_ntt_pass: 64 B per element per pass (read + write), 3 vectors per launch.
    """
    out = None
    try:
        ncu_traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        ncu_traffic = {}
    acc_l, acc_ms = stats.get("k_accumulate<G1>", (0, 0.0))
    if acc_l and acc_ms > 0:
        # exact number of digit entries the G1 accumulate launches consumed (counted by the library in profile mode)
        total_entries = float(stats.get("#msm_entries<G1>", (0, 0.0))[0])
        alg_bytes = total_entries * 68.0
        alg_imad = total_entries * 1360.0
        sec = acc_ms * 1e-3
        traffic = ncu_traffic.get("k_accumulate<G1>", {})
        out = {"kernel": "k_accumulate<G1>", "bound": "int_pipe", "achieved": alg_imad / sec / 1e12, "peak": imad_wide_peak / 1e12,
               "unit": "T IMAD.WIDE/s", "frac": alg_imad / sec / imad_wide_peak if imad_wide_peak else None,
               "peak_source": "measured in this run (ugb_bench_int_pipe, best of three IMAD.WIDE chain shapes)",
               "peak_arithmetic": 148 * 32 * 1.965e9 / 1e12,
               "launches": acc_l, "ms_per_launch": acc_ms / acc_l, "entries_per_launch": total_entries / acc_l,
               "algorithmic": "1360 IMAD.WIDE (10 Montgomery products, madd-2008-s) and 68 B per sorted digit entry",
               "traffic": traffic.get("bytes_per_launch"),
               "traffic_note": "ncu dram bytes of the largest launch (%s)" % traffic.get("launch"),
               "hbm": {"bound": "hbm", "achieved": alg_bytes / sec / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": alg_bytes / sec / 1e9 / hbm_peak,
                       "peak_source": peak_src}}
    ntt = None
    ms = sum(v[1] for k, v in stats.items() if k.startswith("k_ntt_pass"))
    cnt = sum(v[0] for k, v in stats.items() if k.startswith("k_ntt_pass"))
    if cnt and ms > 0:
        alg = cnt * 3 * n * 64.0  # every launch reads and writes 3 vectors of n 32-byte elements
        sec = ms * 1e-3
        ntt = {"kernel": "k_ntt_pass", "bound": "hbm", "achieved": alg / sec / 1e9, "peak": hbm_peak, "unit": "GB/s",
               "frac": alg / sec / 1e9 / hbm_peak, "peak_source": peak_src,
               "traffic": ncu_traffic.get("k_ntt_pass", {}).get("bytes_per_launch"), "launches": cnt, "ms_per_launch": ms / cnt,
               "note": "per pass: 64 B/element moved but log2(n)/(2*passes) Montgomery products of 136 IMAD.WIDE per element"}
    return out, ntt


if __name__ == "__main__":
    main()
