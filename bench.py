#!/usr/bin/env python
"""Benchmark of the commit hot path (BASELINE.json configs[1]): coset LDE + Poseidon2 Merkle commitment of a
synthetic 2^24-row x 256-column BabyBear trace at log_blowup 1, one such trace ("segment") per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b2stark|reference]
                    [--log-height 24] [--width 256] [--log-blowup 1]

Prints ONE JSON line (rank 0). `value` = algorithmic LDE+commit bytes per second over all GPUs, inputs
resident in HBM; `e2e` = the same through the C ABI from pinned host memory (H2D inside the timed region).
`--impl reference` times the CPU oracle (the reference's own CPU prover cannot be built here: no Rust
toolchain, dependencies not vendored -- DESIGN.md) on a bounded sample of the same workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# the reference's stage gauges (crates/prof/data/metrics.example.json): six disjoint stages, then the whole proof
GAUGES = ["main_trace_commit_time_ms", "generate_perm_trace_time_ms", "perm_trace_commit_time_ms",
          "quotient_poly_compute_time_ms", "quotient_poly_commit_time_ms", "pcs_opening_time_ms",
          "stark_prove_excluding_trace_time_ms"]
METRIC = "lde_commit_throughput"
UNIT = "GB/s"


def algorithmic_bytes(log_h, width, b):
    """A_commit = (1 + 2^b) N W 4 + 2^(b+1) N 32   (SURVEY.md 8d / BASELINE.md 3)."""
    n = 1 << log_h
    return (1 + (1 << b)) * n * width * 4 + (2 << b) * n * 32


def permutations(log_h, width, b):
    rows = (1 << log_h) << b
    return rows * ((width + 7) // 8) + rows - 1


def workload_name(log_h, width, b):
    return f"synthetic standalone commit: 2^{log_h}-row x {width}-col BabyBear trace, coset LDE + Poseidon2 Merkle, log_blowup {b}"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nme, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_oracle_sample(log_h, width, b, threads, repeats=1):
    """Time the CPU oracle's commit on a bounded sample (same column count, fewer rows)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding
    orc = oracle_binding.load()
    m = orc.fill_synthetic(1 << log_h, width, 0xB200000000000001)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        orc.commit([m], b, threads=threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best


def run_reference(args):
    """--impl reference: the CPU oracle on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    s_log_h = args.cpu_sample_log_height
    a_bytes = algorithmic_bytes(s_log_h, args.width, args.log_blowup)
    for _ in range(args.warmup if args.warmup < 2 else 1):
        cpu_oracle_sample(s_log_h, args.width, args.log_blowup, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_oracle_sample(s_log_h, args.width, args.log_blowup, threads)
    dt = (time.perf_counter() - t0) / args.steps
    val = a_bytes / dt / 1e9
    sample = f"2^{s_log_h} x {args.width} rows of the same synthetic trace per step (full workload is 2^{args.log_height})"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": workload_name(args.log_height, args.width, args.log_blowup), "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b2stark", choices=["b2stark", "reference"])
    ap.add_argument("--log-height", type=int, default=24)
    ap.add_argument("--width", type=int, default=256)
    ap.add_argument("--log-blowup", type=int, default=1)
    ap.add_argument("--cpu-sample-log-height", type=int, default=16)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-proof", action="store_true")
    ap.add_argument("--sharded", action="store_true",
                    help="N > 1 only: also time ONE commitment of the same trace sharded across the N GPUs (strong scaling; "
                         "openvm_b200/sharded.py, last NTT pass storing into the peers' buffers over NVLink)")
    ap.add_argument("--proof-scale-log", type=int, default=0, help="log2 scale of the shape-replay segment heights")
    ap.add_argument("--proof-engines", type=int, default=8,
                    help="engines proving concurrently on each GPU for the segments/s figure (one host thread each)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from openvm_b200 import B2Engine, FriParameters

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    # run each rank on the CPUs next to its GPU, so that the pinned host buffer of the end-to-end leg is allocated on the
    # GPU's NUMA node (8 ranks streaming 16 GiB each otherwise cross the socket interconnect)
    all_cpus = os.sched_getaffinity(0)
    numa = None
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(local_rank).uuid))
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        numa = sorted(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    log_h, width, b = args.log_height, args.width, args.log_blowup
    H = 1 << log_h
    eng = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(b), local_rank)
    dev = eng.device()
    stream = torch.cuda.ExternalStream(eng.stream, device=local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # one independent trace ("segment") per GPU: weak scaling, no data-path collective
    trace = dev.alloc_matrix(H, width)
    dev.fill_synthetic(trace, 0xB200000000000001 + rank)
    eng.synchronize()

    def step():
        root, pcs = dev.commit([trace])
        pcs.free()
        return root

    for _ in range(args.warmup):
        step()
    eng.reset_metrics()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        root = step()
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    met = eng.metrics()
    if world > 1:
        t = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    a_bytes = algorithmic_bytes(log_h, width, b)
    value = world * a_bytes / (ms_step * 1e-3) / 1e9

    # ---- end to end through the C ABI from pinned host memory
    e2e = None
    nbytes = H * width * 4
    if not args.no_e2e:
        # every rank holds its own host copy of the trace: refuse to drive the box out of memory
        try:
            avail = [int(l.split()[1]) * 1024 for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0]
        except Exception:
            avail = None
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world))
        short = avail is not None and avail < 1.25 * nbytes * local_world
        if world > 1:
            t = torch.tensor([1.0 if short else 0.0], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            short = bool(t.item() > 0)
        if short:
            e2e = {"skipped": f"host memory: {avail} bytes available for {local_world} ranks x {nbytes} bytes"}
    if not args.no_e2e and e2e is None:
        try:
            host = torch.empty(H * width, dtype=torch.int32, pin_memory=True)
        except RuntimeError:
            host = torch.empty(H * width, dtype=torch.int32)
        # host copy of the trace in the reference's in-memory form (column-major DeviceMatrix, Montgomery words)
        dev.download_raw(trace, host.data_ptr(), layout="col", repr="monty")

        def e2e_step():
            # b2s_commit_host: upload of column groups overlapped with LDE + hashing of the previous group; the root
            # comes back to the host inside the call (32 B D2H)
            r, pcs = dev.commit_host(host_ptr=host.data_ptr(), height=H, width=width, layout="col", repr="monty")
            pcs.free()
            return r

        r0 = e2e_step()
        assert list(r0) == list(root), "e2e root differs from device-resident root"
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            e2e_step()
        e1.record(stream)
        barrier()
        ems = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ems], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ems = float(t.item())
        e2e = {"value": world * a_bytes / (ems / args.steps * 1e-3) / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 32, "ms_per_step": ems / args.steps,
               "pinned": bool(host.is_pinned())}
        del host

    # ---- one full STARK proof of a shape-replay RV32IM segment per step, one segment per GPU
    seg = None
    if not args.no_proof:
        from openvm_b200 import AirProvingContext, airs as A
        peng = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(b), local_rank)
        pdev = peng.device()
        shapes = A.segment_shapes(args.proof_scale_log)
        air_objs = [A.segment_air(w) for _, w in shapes]
        key = A.build_key(air_objs, None, log_up_pow_bits=16)
        pk = pdev.transport_pk_to_device(key)
        ctx, cells, perm_cells = [], 0, 0
        for i, (lh, w) in enumerate(shapes):
            parts, pv = A.chain_trace(air_objs[i], 1 << lh, seed=100 * rank + i)
            ctx.append(AirProvingContext(i, pdev.transport_matrix_to_device(parts[-1]), pv))
            cells += w << lh
            perm_cells += (4 * (len(air_objs[i].chunks) + 1)) << lh
        proof = peng.prove(pk, ctx)
        peng.verify(pk, proof)   # the library's own verifier (b2s_verify); raises if the proof is not accepted
        peng.reset_metrics()
        barrier()
        pstream = torch.cuda.ExternalStream(peng.stream, device=local_rank)
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        p0.record(pstream)
        for _ in range(args.steps):
            proof = peng.prove(pk, ctx)
        p1.record(pstream)
        barrier()
        wall = (time.perf_counter() - t0) / args.steps * 1e3
        pms = p0.elapsed_time(p1) / args.steps
        pmet = peng.metrics()
        if world > 1:
            t = torch.tensor([pms, wall], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            pms, wall = float(t[0].item()), float(t[1].item())
        seg = {"workload": f"shape-replay RV32IM segment: {len(shapes)} chain AIRs with LogUp interactions (degree-3 constraints, "
                           f"quotient degree 2), {cells} main cells + {perm_cells} after-challenge cells, FRI log_blowup {b}, "
                           f"{peng.fri_params().num_queries} queries, 16 + 16 PoW bits; proof accepted by b2s_verify",
               "device_ms_per_proof": pms, "wall_ms_per_proof": wall, "segments_per_s": world * 1e3 / wall,
               "proof_bytes": len(proof),
               "stage_ms": {k: pmet[k] / args.steps for k in GAUGES}}
        seg["stage_ms"]["sum_of_six"] = sum(seg["stage_ms"][k] for k in GAUGES[:-1])
        # throughput: K engines on this GPU, one host thread each (the reference's StarkProver also keeps several
        # engines on one GPU, crates/sdk/src/prover/stark.rs:62-75): the host part of one proof (transcript, round
        # trips) overlaps the kernels of the others
        K = max(1, args.proof_engines)
        if K > 1:
            import threading
            extra = []
            for k in range(1, K):
                e2 = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(b), local_rank)
                d2 = e2.device()
                pk2 = d2.transport_pk_to_device(key)
                ctx2 = []
                for i, (lh, w) in enumerate(shapes):
                    parts, pv = A.chain_trace(air_objs[i], 1 << lh, seed=100 * rank + i + 1000 * k)
                    ctx2.append(AirProvingContext(i, d2.transport_matrix_to_device(parts[-1]), pv))
                e2.prove(pk2, ctx2)
                extra.append((e2, pk2, ctx2))
            workers = [(peng, pk, ctx)] + extra
            nper = max(args.steps, 4)

            def work(w):
                for _ in range(nper):
                    w[0].prove(w[1], w[2])

            barrier()
            ths = [threading.Thread(target=work, args=(w,)) for w in workers]
            t0 = time.perf_counter()
            for th in ths:
                th.start()
            for th in ths:
                th.join()
            twall = time.perf_counter() - t0
            if world > 1:
                t = torch.tensor([twall], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                twall = float(t.item())
            seg["segments_per_s_single_engine"] = seg["segments_per_s"]
            seg["segments_per_s"] = world * K * nper / twall
            seg["engines_per_gpu"] = K
            seg["proofs_timed"] = world * K * nper
            for e2, pk2, _ in extra:
                pk2.free()
                e2.close()
        pk.free()
        peng.close()
        if rank == 0 and not args.no_cpu:
            # CPU oracle prover on a bounded sample: the same segment with every height scaled by 2^-5
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_binding
            orc = oracle_binding.load()
            sl = args.proof_scale_log - 5
            sshapes = A.segment_shapes(sl)
            sair = [A.segment_air(w) for _, w in sshapes]
            skey = A.build_key(sair, None, log_up_pow_bits=16)
            sin, scells = [], 0
            for i, (lh, w) in enumerate(sshapes):
                parts, pv = A.chain_trace(sair[i], 1 << lh, seed=i)
                sin.append((i, parts, pv))
                scells += w << lh
            os.sched_setaffinity(0, all_cpus)
            threads = len(all_cpus)
            t0 = time.perf_counter()
            sproof = orc.stark_prove(skey, sin, b, peng.fri_params().num_queries, 16, threads=threads)
            cpu_ms = (time.perf_counter() - t0) * 1e3
            seg["cpu_baseline"] = {"kind": "port", "cores": threads, "ms_per_proof": cpu_ms, "main_cells": scells,
                                   "cells_per_s": scells / cpu_ms * 1e3,
                                   "sample": f"same segment with heights scaled by 2^-5 ({scells} main cells), oracle prover"}
            seg["main_cells_per_s"] = cells * seg["segments_per_s"]

    # ---- opt-in: one commitment of the workload's trace across all ranks (strong scaling)
    sharded_leg = None
    if args.sharded and world > 1 and world & (world - 1) == 0:
        from openvm_b200 import sharded
        trace.free()
        seng = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(b), local_rank)
        sdev = seng.device()
        lo, hi = sharded.column_shards(width, world)[rank]
        shard = sdev.alloc_matrix(H, hi - lo)
        sdev.fill_synthetic(shard, 0xB200000000000001 + rank)
        committer = sharded.ShardedCommitter(seng, H, width, b)
        for _ in range(2):
            sroot, spcs = committer.commit(shard)
            del spcs
        barrier()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for _ in range(args.steps):
            sroot, spcs = committer.commit(shard)
            del spcs
        s1.record()
        barrier()
        t = torch.tensor([s0.elapsed_time(s1) / args.steps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sms = float(t.item())
        sharded_leg = {"what": "ONE commitment of the workload's trace, columns sharded over the ranks, rows exchanged by the last "
                               "NTT pass storing into the peers' buffers over NVLink (no NCCL on the data path)",
                       "scaling": "strong", "ms_per_commit": sms, "value": a_bytes / (sms * 1e-3) / 1e9, "unit": UNIT,
                       "speedup_vs_one_gpu_step": ms_step / sms}
        committer.close()
        seng.close()

    if rank == 0:
        peak, peak_src = measured_peaks()
        steps = args.steps
        rows = H << b
        leaf_ms = met["leaf_hash_ms"] / max(1, met["leaf_hash_launches"])
        leaf_bytes = rows * width * 4 + rows * 32
        lde_bytes = (1 + (1 << b)) * H * width * 4
        lde_ms_step = met["lde_ms"] / steps
        perms = permutations(log_h, width, b)
        roofline = {
            "kernel": "leaf_mixed_kernel (Poseidon2 leaf sponge on the integer-multiply and FP64 pipes side by side; ncu: fmaheavy 78 % + fp64 46 % busy, issue 0.71/cycle, dispatch-bound, dram 1.8 %, profiles/r1g_fp64_pipe.md; reported against HBM as the contract asks)",
            "bound": "hbm", "achieved": leaf_bytes / (leaf_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
            "frac": leaf_bytes / (leaf_ms * 1e-3) / 1e9 / peak,
            # dram__bytes_read.sum + dram__bytes_write.sum of this launch, ncu --set full capture profiles/r1g_fp64_pipe.md
            "traffic": 35.508e9 if (log_h, width, b) == (24, 256, 1) else None, "peak_source": peak_src,
            "ms_per_launch": leaf_ms, "share_of_step": met["leaf_hash_ms"] / steps / ms_step,
        }
        roofline_lde = {
            "kernel": "ntt_strided_kernel / ntt_contig_kernel passes (iNTT + coset NTT)", "bound": "hbm",
            "achieved": lde_bytes / (lde_ms_step * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
            "frac": lde_bytes / (lde_ms_step * 1e-3) / 1e9 / peak, "ms_per_step": lde_ms_step,
            "launches_per_step": met["lde_launches"] / steps, "share_of_step": lde_ms_step / ms_step,
        }
        # integer-pipe view of the hashing: permutations/s vs the register-resident micro-benchmark
        nthreads = 148 * 2048
        pms = eng.bench_poseidon2(nthreads, 64)
        perm_peak = nthreads * 64 / (pms * 1e-3)
        mms = eng.bench_modmul(nthreads, 4096, 8)
        int_pipe = {
            "poseidon2_perm_per_s": perms / ((met["leaf_hash_ms"] + met["compress_ms"]) / steps * 1e-3),
            "poseidon2_perm_per_s_register_peak": perm_peak,
            "modmul_per_s_peak": nthreads * 4096 * 8 / (mms * 1e-3),
        }
        # the bound that explains the time (profiles/r1j_dispatch_probe.md): S-box products of the hashing
        # (564 per permutation) and butterflies of the LDE (one product each) against the measured product rate
        int_pipe["hash_products_per_s"] = int_pipe["poseidon2_perm_per_s"] * 564
        int_pipe["hash_frac_of_modmul_peak"] = int_pipe["hash_products_per_s"] / int_pipe["modmul_per_s_peak"]
        butterflies = (1 + (1 << b)) * 0.5 * H * log_h * width
        int_pipe["lde_butterflies_per_s"] = butterflies / (lde_ms_step * 1e-3)
        int_pipe["lde_frac_of_modmul_peak"] = int_pipe["lde_butterflies_per_s"] / int_pipe["modmul_per_s_peak"]
        cpu = None
        if not args.no_cpu:
            os.sched_setaffinity(0, all_cpus)
            threads = len(all_cpus)
            s = args.cpu_sample_log_height
            dt = cpu_oracle_sample(s, width, b, threads)
            cpu = {"value": algorithmic_bytes(s, width, b) / dt / 1e9, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": f"2^{s} x {width} rows of the same synthetic trace, one commit, {dt:.2f} s"}
        launches = met["lde_launches"] + met["leaf_hash_launches"] + met["compress_launches"] + met["other_launches"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic",
            "config": {"workload": workload_name(log_h, width, b), "segments_per_gpu": 1,
                       "cpu_affinity": f"rank 0 pinned to {len(numa)} CPUs next to its GPU" if numa else "unpinned",
                       "l2": "inputs (16 GiB trace, 32 GiB LDE) are far larger than the 126 MB L2; no flush needed",
                       "fri": {"log_blowup": b}},
            "roofline": roofline, "roofline_lde": roofline_lde, "int_pipe": int_pipe, "cpu_baseline": cpu,
            "sharded_commit": sharded_leg,
            "e2e": e2e, "segment_proof": seg, "gpu_launches": int(launches), "clocks": clocks,
            "stage_ms_per_step": {"main_trace_commit_time_ms": met["commit_total_ms"] / steps,
                                  "lde_ms": lde_ms_step, "leaf_hash_ms": met["leaf_hash_ms"] / steps,
                                  "compress_ms": met["compress_ms"] / steps},
            "root": [int(x) for x in root],
        }
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    trace.free()
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
