#!/usr/bin/env python
"""Benchmark of the STARK proving hot path on B200 (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b2stark|reference] [--only LEGS] [--skip LEGS]

Headline (configs[1]): coset LDE + Poseidon2 Merkle commitment of a synthetic 2^24-row x 256-column BabyBear trace at
log_blowup 1, one such trace per GPU. `value` = algorithmic LDE+commit bytes per second over all GPUs with the trace
resident in HBM; `e2e` = the same through the C ABI from pinned host memory (H2D inside the timed region).
Further legs in the same JSON line (each time-boxed, each checked by the ORACLE outside its timed region):
  commit_b2      the same commit at log_blowup 2
  segment        one STARK proof of the shape-replay RV32IM segment (the reference's 34-AIR example table), stage gauges,
                 K engines per GPU for segments/s
  keccak / leaf  the KeccakVmAir shape (config 4) and the leaf-aggregation shape at log_blowup 2 (config 5)
  dag            S segments -> S leaves -> arity-3 internal tree as ONE FIFO over the engines of all ranks (configs 3, 5)
  sharded        (N > 1) ONE commitment of the workload's trace across the N GPUs, rows exchanged over NVLink by the NTT
`--impl reference` times the CPU arm: oracle/fast_commit.cpp (AVX2, Montgomery, all host threads -- the reference's own
CPU prover is Rust with un-vendored dependencies and cannot be built here) on a bounded sample of the same
workload.
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
# Several engines per GPU, each with a dozen streams: with the default 8 hardware queues unrelated streams share a queue and
# wait on each other's latency-bound kernels (34-AIR segment, 8 engines: 59 -> 64 segments/s with 32). Must be set before CUDA
# initialises; a deployment sets it the same way.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

# the reference's stage gauges (crates/prof/data/metrics.example.json): six disjoint stages, then the whole proof
GAUGES = ["main_trace_commit_time_ms", "generate_perm_trace_time_ms", "perm_trace_commit_time_ms",
          "quotient_poly_compute_time_ms", "quotient_poly_commit_time_ms", "pcs_opening_time_ms",
          "stark_prove_excluding_trace_time_ms"]
OPENING_PARTS = ["ood_eval_ms", "reduced_opening_ms", "fri_commit_phase_ms"]
METRIC = "lde_commit_throughput"
UNIT = "GB/s"
SEED = 0xB200000000000001
ALL_LEGS = ["commit", "e2e", "commit_b2", "segment", "keccak", "leaf", "dag", "sharded", "cpu"]
# the reference's own example run (crates/prof/data/metrics.example.json, unknown CPU, log_blowup 2): context, not a baseline
REFERENCE_EXAMPLE = {"source": "crates/prof/data/metrics.example.json (hardware not stated)",
                     "ecrecover_segment": {"cells": 59236298, "stark_prove_excluding_trace_time_ms": 2391,
                                           "main_trace_commit_time_ms": 417, "main_cells_committed": 15204352 + 14680064},
                     "leaf": {"main_cells": 445.7e6, "stark_prove_excluding_trace_time_ms": 34900,
                              "main_trace_commit_time_ms": 6309}}


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


def oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding
    return oracle_binding.load()


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_commit(orc, log_h, width, b, threads, seed=SEED):
    """The CPU arm's commit of the synthetic trace (oracle/fast_commit.cpp). Returns (root, seconds WITHOUT trace generation)."""
    root, t = orc.fast_commit(log_h, width, b, seed=seed, threads=threads)
    return root, t["lde_s"] + t["hash_s"], t


# ---- --impl reference ---------------------------------------------------------------------------------------------------
def run_reference(args):
    """The CPU implementation of the path on the host cores, rank 0 only. Every step commits a bounded sample (fewer rows,
    same columns, same generator) sized from a probe so that the whole run ends within a few minutes; the trace is
    generated outside the timed figure (the GPU arm's trace is resident before its timing starts)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    orc = oracle()
    threads = host_threads()
    w, b = args.width, args.log_blowup
    probe_log = min(args.log_height, 18)
    _, probe_s, _ = cpu_commit(orc, probe_log, w, b, threads)
    budget = args.reference_budget_s / max(1, args.steps + min(args.warmup, 1))
    s_log = probe_log
    while s_log < args.log_height and probe_s * (1 << (s_log + 1 - probe_log)) * 1.15 <= budget:
        s_log += 1
    if args.cpu_sample_log_height:
        s_log = args.cpu_sample_log_height
    for _ in range(min(args.warmup, 1)):
        cpu_commit(orc, s_log, w, b, threads)
    secs = []
    for _ in range(args.steps):
        secs.append(cpu_commit(orc, s_log, w, b, threads)[1])
    dt = sum(secs) / len(secs)
    val = algorithmic_bytes(s_log, w, b) / dt / 1e9
    sample = (f"2^{s_log} x {w} of the same synthetic trace per step" +
              (" (the full workload)" if s_log == args.log_height else f" (full workload: 2^{args.log_height} rows)") +
              ", AVX2 Montgomery port of the commit, trace generation excluded")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": workload_name(args.log_height, w, b), "sample": sample,
                   "same_config": s_log == args.log_height},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "reference_example": REFERENCE_EXAMPLE,
    }
    print(json.dumps(line), flush=True)


# ---- helpers of the GPU arm ---------------------------------------------------------------------------------------------
class Ctx:
    """Per-process state shared by the legs."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.args = torch, dist, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        if os.environ.get("B2S_BENCH_BACKEND") == "none":   # diagnostic: ranks run as unrelated processes
            self.world = 1
        self.all_cpus = sorted(os.sched_getaffinity(0))
        self.cpus = self.all_cpus
        self.affinity = "unpinned"
        self.reduce_device = "cuda"
        self.sched = "default"

    def pin_cpus(self):
        """Each rank gets a DISJOINT share of the CPUs next to its GPU (NUMA-local pinned buffers; 8 ranks x K proving threads
        no longer compete for the same cores)."""
        torch = self.torch
        near = self.all_cpus
        if os.environ.get("B2S_BENCH_PIN", "1") == "0":
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(self.local_rank).uuid))
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.local_rank)
            pynvml.nvmlDeviceSetCpuAffinity(h)
            near = sorted(os.sched_getaffinity(0))
        except Exception:
            pass
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", self.world))
        share = near
        if local_world > 1 and len(near) >= 2 * local_world:
            # ranks that see the same CPU set split it evenly
            per = len(near) // local_world
            share = near[self.local_rank * per:(self.local_rank + 1) * per]
        try:
            os.sched_setaffinity(0, share)
            self.cpus = share
            self.affinity = f"{len(share)} of {len(self.all_cpus)} CPUs ({share[0]}-{share[-1]}), disjoint per rank"
        except OSError:
            pass

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def allmax(self, *vals):
        if self.world == 1:
            return vals if len(vals) > 1 else vals[0]
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.reduce_device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        out = [float(x) for x in t.tolist()]
        return out if len(out) > 1 else out[0]

    def engine(self, b):
        from openvm_b200 import B2Engine, FriParameters
        return B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(b), self.local_rank)


def trim_pool(cx):
    """Between legs: give the device's stream-ordered pool back to the driver. Engines of the previous leg are closed, but the
    pool (release threshold = never) keeps their blocks; after a 12-engine leg its free list made every later allocation
    slower (FRI commit phase of the next single-engine proof 3.3 -> 7-14 ms: ~400 small allocations per proof)."""
    try:
        import ctypes
        rt = ctypes.CDLL("libcudart.so.12")
        cx.torch.cuda.synchronize()
        pool = ctypes.c_void_p()
        if rt.cudaDeviceGetDefaultMemPool(ctypes.byref(pool), cx.local_rank) == 0:
            rt.cudaMemPoolTrimTo(pool, ctypes.c_size_t(0))
    except OSError:
        pass


def leg_commit(cx, b, steps, warmup, sample_clocks):
    """Device-resident commit of the workload's trace: K timed steps between CUDA events on the engine's stream."""
    torch, args = cx.torch, cx.args
    H, width = 1 << args.log_height, args.width
    eng = cx.engine(b)
    dev = eng.device()
    stream = torch.cuda.ExternalStream(eng.stream, device=cx.local_rank)
    trace = dev.alloc_matrix(H, width)
    dev.fill_synthetic(trace, SEED + cx.rank)
    eng.synchronize()
    root = None

    def step():
        r, pcs = dev.commit([trace])
        pcs.free()
        return r

    for _ in range(warmup):
        step()
    eng.reset_metrics()
    sampler = ClockSampler(cx.local_rank) if sample_clocks else None
    cx.barrier()
    if sampler:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(steps):
        root = step()
    ev1.record(stream)
    cx.barrier()
    clocks = sampler.stop() if sampler else None
    ms_step = cx.allmax(ev0.elapsed_time(ev1)) / steps
    met = eng.metrics()
    return {"eng": eng, "dev": dev, "trace": trace, "stream": stream, "ms_step": ms_step, "met": met, "root": root,
            "clocks": clocks, "steps": steps, "b": b}


def leg_e2e(cx, base, steps):
    """The same commit through b2s_commit_host from pinned host memory; plus the host's H2D ceiling with all ranks copying."""
    torch, args = cx.torch, cx.args
    H, width, b = 1 << args.log_height, args.width, base["b"]
    nbytes = H * width * 4
    try:
        avail = [int(l.split()[1]) * 1024 for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0]
    except Exception:
        avail = None
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", cx.world))
    short = avail is not None and avail < 1.25 * nbytes * local_world
    if cx.allmax(1.0 if short else 0.0) > 0:
        return {"skipped": f"host memory: {avail} bytes available for {local_world} ranks x {nbytes} bytes"}
    dev, trace, stream = base["dev"], base["trace"], base["stream"]
    try:
        host = torch.empty(H * width, dtype=torch.int32, pin_memory=True)   # allocated after pin_cpus(): NUMA-local
    except RuntimeError:
        host = torch.empty(H * width, dtype=torch.int32)
    # host copy of the trace in the reference's in-memory form (column-major DeviceMatrix, Montgomery words)
    dev.download_raw(trace, host.data_ptr(), layout="col", repr="monty")

    def e2e_step():
        # upload of column groups overlapped with LDE + hashing of the previous group; the root comes back inside the call
        r, pcs = dev.commit_host(host_ptr=host.data_ptr(), height=H, width=width, layout="col", repr="monty")
        pcs.free()
        return r

    r0 = e2e_step()
    assert list(r0) == list(base["root"]), "e2e root differs from device-resident root"
    cx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        e2e_step()
    e1.record(stream)
    cx.barrier()
    ems = cx.allmax(e0.elapsed_time(e1)) / steps
    a_bytes = algorithmic_bytes(args.log_height, width, b)
    out = {"value": cx.world * a_bytes / (ems * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": nbytes,
           "d2h_bytes_per_step": 32, "ms_per_step": ems, "pinned": bool(host.is_pinned())}
    # H2D ceiling of this host: every rank copies 4 GiB from its pinned buffer at the same time, nothing else running
    chunk = min(nbytes, 4 << 30)
    dst = torch.empty(chunk // 4, dtype=torch.int32, device="cuda")
    src = host[:chunk // 4]
    dst.copy_(src, non_blocking=True)
    cx.barrier()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    for _ in range(2):
        dst.copy_(src, non_blocking=True)
    c1.record()
    cx.barrier()
    cms = cx.allmax(c0.elapsed_time(c1)) / 2
    per_gpu = chunk / (cms * 1e-3) / 1e9
    out["h2d_ceiling_gbs_per_gpu"] = per_gpu
    out["h2d_ceiling_gbs"] = per_gpu * cx.world
    out["h2d_floor_ms_per_step"] = nbytes / per_gpu / 1e6
    out["note"] = ("the upload of a step cannot take less than h2d_floor_ms_per_step on this host with all ranks copying; "
                   "the deployment path is zero-copy adoption of device traces (segment / dag legs), not bigger host copies")
    del dst, host
    return out


def dag_weights(air):
    """M_Q per row of the quotient domain: sum over DAG nodes of 1 (F x F), 4 (EF x F) or 16 (EF x EF) for
    products, 0 otherwise, plus 16 per constraint (the alpha fold)."""
    from openvm_b200 import airs as A
    ext = []
    total = 0
    for kind, a, b_, _c in air.dag.nodes:
        if kind in (A.N_PERM, A.N_CHALLENGE, A.N_EXPOSED):
            e = True
        elif kind in (A.N_ADD, A.N_SUB, A.N_MUL):
            e = ext[a] or ext[b_]
        elif kind == A.N_NEG:
            e = ext[a]
        else:
            e = False
        if kind == A.N_MUL:
            total += 16 if (ext[a] and ext[b_]) else (4 if e else 1)
        ext.append(e)
    return total + 16 * len(air.dag.constraints)


class ShapeProver:
    """One engine proving one multi-AIR shape: key built from authored chain AIRs, traces generated ON THE DEVICE by a kernel on
    the engine's stream, which is a caller-owned (torch) stream shared through b2s_engine_set_stream -- the reference's device
    tracegen -> prover hand-off (crates/vm/src/system/cuda/program.rs:60,119-125); the adoption of caller-owned buffers
    (b2s_matrix_adopt_device) is exercised by tests/test_gpu_prove.py."""

    def __init__(self, cx, shapes, b, log_up_pow_bits=16, key=None, air_objs=None):
        from openvm_b200 import airs as A
        torch = cx.torch
        self.cx, self.shapes, self.b = cx, shapes, b
        self.air_objs = air_objs or [A.segment_air(w, degree=dg, perm_cols=pc) for _, w, pc, dg in shapes]
        self.key = key if key is not None else A.build_key(self.air_objs, None, log_up_pow_bits=log_up_pow_bits)
        self.eng = cx.engine(b)
        self.dev = self.eng.device()
        self.tstream = torch.cuda.Stream(device=cx.local_rank)
        self.eng.set_stream(self.tstream.cuda_stream)
        t0 = time.perf_counter()
        self.pk = self.dev.transport_pk_to_device(self.key)
        self.key_load_s = time.perf_counter() - t0
        self.cells = sum(w << lh for lh, w, _, _ in shapes)
        self.perm_cells = sum((4 * (len(a.chunks) + 1)) << lh for a, (lh, _, _, _) in zip(self.air_objs, shapes))
        self.ctx, self.mats = None, None

    def generate(self, seed):
        """Traces of one proof, written by a device kernel on the engine's stream (b2s_matrix_fill_chain, the stand-in for a
        device tracegen): one launch per AIR, no host copy, no synchronisation."""
        from openvm_b200 import AirProvingContext
        if self.mats is None:
            self.mats = [self.dev.alloc_matrix(1 << lh, w) for lh, w, _, _ in self.shapes]
        ctx = []
        for i, m in enumerate(self.mats):
            ctx.append(AirProvingContext(i, m, self.dev.fill_chain(m, seed * 1000 + i)))
        self.ctx = ctx
        return ctx

    def prove(self, seed=None):
        if seed is not None or self.ctx is None:
            self.generate(seed or 0)
        with self.cx.torch.cuda.stream(self.tstream):
            return self.eng.prove(self.pk, self.ctx)

    def m_q(self):
        return sum(dag_weights(a) << (lh + a.log_qd) for a, (lh, _, _, _) in zip(self.air_objs, self.shapes))

    def opened_lde_bytes(self):
        """bytes of every committed LDE column the reduced openings read once: main + after-challenge + quotient chunks"""
        tot = 0
        for a, (lh, w, _, _) in zip(self.air_objs, self.shapes):
            tot += ((w + 4 * (len(a.chunks) + 1) + (4 << a.log_qd)) * 4) << (lh + self.b)
        return tot

    def close(self):
        self.ctx = self.mats = None
        self.pk.free()
        self.eng.close()


def leg_shape(cx, name, shapes, b, steps, what, orc=None, modmul_peak=None, hbm_peak=None, cpu_scale_log=None):
    """One proof per step of a multi-AIR shape; stage gauges; acceptance by the product's verifier and by the ORACLE's."""
    torch = cx.torch
    sp = ShapeProver(cx, shapes, b)
    sp.generate(100 + cx.rank)
    proof = sp.prove()
    sp.eng.verify(sp.pk, proof)          # b2s_verify: raises if not accepted
    accepted = ["b2s_verify"]
    q = sp.eng.fri_params().num_queries
    if orc is not None and cx.rank == 0:
        res = orc.stark_verify(sp.key, proof, b, q, 16)
        assert res is None, f"{name}: the oracle verifier rejects the device proof: {res}"
        accepted.append("oracle verifier (oracle/stark_verify.hpp)")
    sp.eng.reset_metrics()
    cx.barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    p0.record(sp.tstream)
    for _ in range(steps):
        proof = sp.prove()
    p1.record(sp.tstream)
    sp.tstream.synchronize()
    wall = (time.perf_counter() - t0) / steps * 1e3
    cx.barrier()
    pms, wall = cx.allmax(p0.elapsed_time(p1) / steps, wall)
    m = sp.eng.metrics()
    stage = {k: m[k] / steps for k in GAUGES}
    stage["sum_of_six"] = sum(stage[k] for k in GAUGES[:-1])
    out = {"workload": f"{what}: {len(shapes)} AIRs, {sp.cells} main cells + {sp.perm_cells} after-challenge cells, FRI log_blowup "
                       f"{b}, {q} queries, 16 + 16 PoW bits; traces generated on the device, no host copy",
           "accepted_by": accepted, "device_ms_per_proof": pms, "wall_ms_per_proof": wall, "proofs_per_s": cx.world * 1e3 / wall,
           "proof_bytes": len(proof), "stage_ms": stage, "opening_parts_ms": {k: m[k] / steps for k in OPENING_PARTS},
           "key_load_s": sp.key_load_s, "device_peak_gib": sp.eng.device_memory()[1] / 2 ** 30,
           "quotient_kernels": {"generated": m["quotient_jit_launches"] // steps,
                                "interpreted": m["quotient_interpreter_launches"] // steps},
           "cells_per_s": cx.world * (sp.cells + sp.perm_cells) / wall * 1e3}
    if modmul_peak:
        mq = sp.m_q()
        qms = stage["quotient_poly_compute_time_ms"]
        out["roofline_quotient"] = {"bound": "integer pipe (modular products)", "M_Q": mq, "achieved": mq / (qms * 1e-3),
                                    "peak": modmul_peak, "unit": "modmul/s", "frac": mq / (qms * 1e-3) / modmul_peak,
                                    "ms": qms}
        rb = sp.opened_lde_bytes()
        rms = out["opening_parts_ms"]["reduced_opening_ms"]
        if rms > 0:
            out["roofline_reduced_opening"] = {"bound": "hbm", "bytes": rb, "achieved": rb / (rms * 1e-3) / 1e9, "peak": hbm_peak,
                                               "unit": "GB/s", "frac": rb / (rms * 1e-3) / 1e9 / hbm_peak, "ms": rms}
    if cpu_scale_log is not None and orc is not None and cx.rank == 0:
        # CPU oracle prover on a bounded sample: the same shape with every height scaled by 2^cpu_scale_log
        from openvm_b200 import airs as A
        sshapes = [(max(1, lh + cpu_scale_log), w, pc, dg) for lh, w, pc, dg in shapes]
        sin, scells = [], 0
        for i, (a, (lh, w, _, _)) in enumerate(zip(sp.air_objs, sshapes)):
            parts, pv = A.chain_trace(a, 1 << lh, seed=i, hashed=True)
            sin.append((i, parts, pv))
            scells += (w + 4 * (len(a.chunks) + 1)) << lh
        threads = len(cx.all_cpus)
        os.sched_setaffinity(0, cx.all_cpus)
        t0 = time.perf_counter()
        orc.stark_prove(sp.key, sin, b, q, 16, threads=threads)
        cpu_ms = (time.perf_counter() - t0) * 1e3
        os.sched_setaffinity(0, cx.cpus)
        out["cpu_baseline"] = {"kind": "port", "cores": threads, "ms_per_proof": cpu_ms, "cells": scells,
                               "cells_per_s": scells / cpu_ms * 1e3,
                               "sample": f"same shape with heights scaled by 2^{cpu_scale_log} ({scells} cells), plain oracle prover"}
    return out, sp


def leg_segment_throughput(cx, sp0, shapes, b, K, steps):
    """K engines on this GPU, one host thread each (the reference's StarkProver also keeps several engines on one GPU,
    crates/sdk/src/prover/stark.rs:62-75): the host part of one proof overlaps the kernels of the others."""
    workers = [sp0]
    for k in range(1, K):
        sp = ShapeProver(cx, shapes, b, key=sp0.key, air_objs=sp0.air_objs)
        sp.generate(1000 * k + cx.rank)
        sp.prove()
        workers.append(sp)
    # every engine proves back to back for a fixed wall time (short fixed counts were +-25 % from run to run: the figure
    # depends on how the engines' host phases interleave, and start-up and tail effects dominated)
    duration = max(2.0, 0.25 * steps)
    counts = [0] * len(workers)
    t_end = [0.0]

    def work(k, w):
        # a new host thread starts on device 0: without this, every rank's engine threads open a context on GPU 0
        cx.torch.cuda.set_device(cx.local_rank)
        while time.perf_counter() < t_end[0]:
            w.prove()
            w.tstream.synchronize()
            counts[k] += 1

    cx.barrier()
    ths = [threading.Thread(target=work, args=(k, w)) for k, w in enumerate(workers)]
    t0 = time.perf_counter()
    t_end[0] = t0 + duration
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    mine = sum(counts) / (time.perf_counter() - t0)
    total = mine
    if cx.world > 1:
        t = cx.torch.tensor([mine], dtype=cx.torch.float64, device=cx.reduce_device)
        cx.dist.all_reduce(t, op=cx.dist.ReduceOp.SUM)
        total = float(t.item())
    slowest = -cx.allmax(-mine)
    for w in workers[1:]:
        w.close()
    return {"segments_per_s": total, "segments_per_s_slowest_gpu": slowest, "engines_per_gpu": K,
            "proofs_timed": int(round(total * duration)), "timed_s": duration}


def leg_dag(cx, S, K, seg_shapes, agg_shapes, orc):
    """The continuation as a DAG (configs 3 and 5): S distinct segments -> S leaf proofs -> arity-3 internal tree
    (crates/sdk/src/prover/agg.rs:124-202, crates/sdk/src/config/mod.rs:26-27; the sequential loop being replaced:
    crates/vm/src/arch/vm.rs:979-1007), ONE FIFO over the engines of all ranks (job -> first free engine,
    openvm_b200/scheduler.py DagFifo over the process group's store; no data-path collective). A parent consumes its children's
    proofs as the hint stream its verifier program would be fed (b2s_proof_hint_stream), then proves its own shape."""
    from openvm_b200 import scheduler
    dist = cx.dist
    kinds, deps = scheduler.aggregation_dag(S)
    if cx.world > 1:
        try:
            store = dist.distributed_c10d._get_default_store()
        except Exception:   # private accessor gone: a store of our own next to the rendezvous port
            store = dist.TCPStore(os.environ.get("MASTER_ADDR", "127.0.0.1"), int(os.environ.get("MASTER_PORT", "29500")) + 17,
                                  cx.world, cx.rank == 0)
    else:
        store = dist.HashStore()
    fifo = scheduler.DagFifo(store, deps, tag=f"dag{S}")
    workers = []
    for k in range(K):
        seg = ShapeProver(cx, seg_shapes, 1, key=workers[0][0].key if workers else None,
                          air_objs=workers[0][0].air_objs if workers else None)
        agg = ShapeProver(cx, agg_shapes, 2, key=workers[0][1].key if workers else None,
                          air_objs=workers[0][1].air_objs if workers else None)
        seg.prove(1); agg.prove(1)     # warm both paths
        workers.append((seg, agg))
    logs = [[] for _ in range(K)]

    def run(k):
        cx.torch.cuda.set_device(cx.local_rank)   # threads start on device 0 (see leg_segment_throughput)
        seg, agg = workers[k]

        def job(jid, children):
            kind, _idx, _layer = kinds[jid]
            sp = seg if kind == "segment" else agg
            for c in children:
                sp.eng.proof_hint_stream(c)            # host work of a parent: its children's proofs as verifier input
            proof = sp.prove(seed=7 + jid)             # distinct traces per job, generated on the device
            sp.tstream.synchronize()
            return proof

        logs[k] = fifo.work(job)

    cx.barrier()
    ths = [threading.Thread(target=run, args=(k,)) for k in range(K)]
    t0 = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    cx.barrier()
    wall = cx.allmax(time.perf_counter() - t0)
    # job costs (ready -> done) from every rank, for the critical path
    mine = [(jid, t_done - t_ready) for lg in logs for jid, _, t_ready, t_done in lg]
    if cx.world > 1:
        gathered = [None] * cx.world
        dist.all_gather_object(gathered, mine)
    else:
        gathered = [mine]
    cost = [0.0] * len(deps)
    ran_on = [len(lst) for lst in gathered]
    for lst in gathered:
        for jid, c in lst:
            cost[jid] = c
    out = None
    if cx.rank == 0:
        root_proof = fifo.result(len(deps) - 1)
        accepted = ["b2s_verify"]
        workers[0][1].eng.verify(workers[0][1].pk, root_proof)
        if orc is not None:
            res = orc.stark_verify(workers[0][1].key, root_proof, 2, workers[0][1].eng.fri_params().num_queries, 16)
            assert res is None, f"dag: oracle verifier rejects the root of the aggregation tree: {res}"
            res = orc.stark_verify(workers[0][0].key, fifo.result(0), 1, workers[0][0].eng.fri_params().num_queries, 16)
            assert res is None, f"dag: oracle verifier rejects segment 0: {res}"
            accepted.append("oracle verifier (root of the tree and segment 0)")
        crit = scheduler.critical_path(deps, cost)
        total = sum(cost)
        engines = cx.world * K
        by_kind = {}
        for (kind, _, _), c in zip(kinds, cost):
            by_kind.setdefault(kind, []).append(c)
        lower = max(crit, total / engines)
        out = {"what": f"{S} distinct segments (34-AIR shape, b = 1) -> {S} leaves -> arity-3 internal tree "
                       f"({len(deps) - 2 * S} internal nodes; aggregation shape: {len(agg_shapes)} AIRs, "
                       f"{workers[0][1].cells} main cells, b = 2), one FIFO over {engines} engines on {cx.world} GPU(s); "
                       "traces generated on the device per job",
               "jobs": len(deps), "wall_s": wall, "jobs_per_s": len(deps) / wall, "segments_per_s": S / wall,
               "critical_path_s": crit, "sum_of_job_s": total, "lower_bound_s": lower,
               "efficiency_vs_lower_bound": lower / wall,
               "jobs_per_rank": ran_on, "mean_job_ms": {k: 1e3 * sum(v) / len(v) for k, v in by_kind.items()},
               "accepted_by": accepted}
    for seg, agg in workers:
        seg.close(); agg.close()
    return out


def leg_sharded(cx, base_ms_step, b, steps):
    """ONE commitment of the workload's trace across all ranks (strong scaling): columns sharded, the last
    NTT pass stores every LDE row into the buffer of the rank that hashes it, over NVLink peer mappings."""
    from openvm_b200 import sharded
    torch, args = cx.torch, cx.args
    H, width = 1 << args.log_height, args.width
    seng = cx.engine(b)
    sdev = seng.device()
    lo, hi = sharded.column_shards(width, cx.world)[cx.rank]
    shard = sdev.alloc_matrix(H, hi - lo)
    sdev.fill_synthetic(shard, SEED + cx.rank)
    committer = sharded.ShardedCommitter(seng, H, width, b)
    for _ in range(2):
        sroot, spcs = committer.commit(shard)
        del spcs
    cx.barrier()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(steps):
        sroot, spcs = committer.commit(shard)
        del spcs
    s1.record()
    cx.barrier()
    sms = cx.allmax(s0.elapsed_time(s1) / steps)
    a_bytes = algorithmic_bytes(args.log_height, width, b)
    lde_bytes_rank = (H << b) * (hi - lo) * 4
    out = {"what": "ONE commitment of the workload's trace, columns sharded over the ranks, rows exchanged by the last NTT pass "
                   "storing into the peers' buffers over NVLink (fused; no NCCL on the data path)",
           "scaling": "strong", "ms_per_commit": sms, "value": a_bytes / (sms * 1e-3) / 1e9, "unit": UNIT,
           "speedup_vs_one_gpu_step": base_ms_step / sms,
           "nvlink_bytes_stored_per_rank": lde_bytes_rank * (cx.world - 1) // cx.world,
           "nvlink_gbs_per_rank_over_the_commit": lde_bytes_rank * (cx.world - 1) / cx.world / (sms * 1e-3) / 1e9}
    committer.close()
    shard.free()
    seng.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b2stark", choices=["b2stark", "reference"])
    ap.add_argument("--log-height", type=int, default=24)
    ap.add_argument("--width", type=int, default=256)
    ap.add_argument("--log-blowup", type=int, default=1)
    ap.add_argument("--cpu-sample-log-height", type=int, default=0, help="reference arm: force the sample's log height")
    ap.add_argument("--reference-budget-s", type=float, default=150.0)
    ap.add_argument("--cpu-full-budget-s", type=float, default=100.0,
                    help="rank 0 runs the CPU arm at the FULL workload size (root check + cpu_baseline) when a probe says it "
                         "fits this budget; otherwise at the largest size that does")
    ap.add_argument("--only", default="", help="comma-separated legs to run (default: all): " + ",".join(ALL_LEGS))
    ap.add_argument("--skip", default="", help="comma-separated legs to skip")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-proof", action="store_true", help="skip segment, keccak, leaf and dag")
    ap.add_argument("--segment-scale-log", type=int, default=0,
                    help="log2 scale of the segment's heights (0 = the reference's example table; 3 = ALU chips of 2^20 rows)")
    ap.add_argument("--leaf-scale-log", type=int, default=0)
    ap.add_argument("--proof-engines", type=int, default=12,
                    help="engines proving concurrently on each GPU for the segments/s figure (one host thread each)")
    ap.add_argument("--dag-segments", type=int, default=16)
    ap.add_argument("--dag-engines", type=int, default=4)
    ap.add_argument("--dag-agg-scale-log", type=int, default=-3)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    legs = set(args.only.split(",")) if args.only else set(ALL_LEGS)
    legs -= set(args.skip.split(",")) if args.skip else set()
    if args.no_e2e:
        legs.discard("e2e")
    if args.no_cpu:
        legs.discard("cpu")
    if args.no_proof:
        legs -= {"segment", "keccak", "leaf", "dag"}
    legs.add("commit")

    import torch
    import torch.distributed as dist
    from openvm_b200 import airs as A

    cx = Ctx(args)
    rank, world = cx.rank, cx.world
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    cx.pin_cpus()
    # K engine threads per GPU spend most of their time waiting for the device. With the default policy they spin, and once the
    # host has about as many waiting threads as CPUs (4 ranks x 8-12 engines on 32 CPUs) the spinners starve the threads that
    # have work: one rank fell to 1.6 proofs/s and every host round trip of the others doubled. cudaDeviceScheduleYield lets a
    # waiting thread give its slice away (1 GPU, 12 engines on 8 CPUs: 67 segments/s against 65 spinning; blocking waits: 58).
    sched = os.environ.get("B2S_BENCH_SCHED", "yield")
    try:
        import ctypes
        rt = ctypes.CDLL("libcudart.so.12")
        flag = {"spin": 1, "yield": 2, "block": 4}.get(sched, 0)
        if flag:
            rt.cudaSetDevice(cx.local_rank)
            rc = rt.cudaSetDeviceFlags(flag)
            cx.sched = f"{sched} (cudaSetDeviceFlags rc {rc})"
    except OSError:
        cx.sched = "default (libcudart not found)"
    torch.cuda.set_device(cx.local_rank)
    if world > 1:
        if os.environ.get("B2S_BENCH_BACKEND", "nccl") == "gloo":   # diagnostic only
            dist.init_process_group("gloo")
            cx.reduce_device = "cpu"
        else:
            dist.init_process_group("nccl", device_id=torch.device("cuda", cx.local_rank))
    orc = oracle() if (rank == 0 and "cpu" in legs) else None
    log_h, width, b = args.log_height, args.width, args.log_blowup
    t_start = time.perf_counter()

    def log(msg):
        if rank == 0:
            sys.stderr.write(f"[bench {time.perf_counter() - t_start:7.1f}s] {msg}\n")
            sys.stderr.flush()

    # ---- headline: device-resident commit, then end to end
    base = leg_commit(cx, b, args.steps, args.warmup, sample_clocks=True)
    ms_step, met, root = base["ms_step"], base["met"], base["root"]
    a_bytes = algorithmic_bytes(log_h, width, b)
    value = world * a_bytes / (ms_step * 1e-3) / 1e9
    log(f"commit b={b}: {ms_step:.1f} ms/step, {value:.1f} GB/s")
    e2e = leg_e2e(cx, base, args.steps) if "e2e" in legs else None
    if e2e and "value" in e2e:
        log(f"e2e: {e2e['ms_per_step']:.1f} ms/step; h2d ceiling {e2e['h2d_ceiling_gbs_per_gpu']:.1f} GB/s per GPU")
    # micro-benchmarks for the integer-pipe view (one modular-product peak for every fraction)
    eng = base["eng"]
    sms = torch.cuda.get_device_properties(cx.local_rank).multi_processor_count
    nthreads = sms * 2048
    pms = eng.bench_poseidon2(nthreads, 64)
    mms = eng.bench_modmul(nthreads, 4096, 8)
    modmul_peak = nthreads * 4096 * 8 / (mms * 1e-3)
    # root of a 2^20-row commit of the same generator (other NTT plan), for the oracle check below
    small_log = min(20, log_h)
    small = base["dev"].alloc_matrix(1 << small_log, width)
    base["dev"].fill_synthetic(small, SEED)
    small_root, spcs = base["dev"].commit([small])
    spcs.free(); small.free()
    base["trace"].free()
    eng.close()
    trim_pool(cx)
    hbm_peak, peak_src = measured_peaks()

    commit_b2 = None
    if "commit_b2" in legs and b != 2:
        r2 = leg_commit(cx, 2, max(2, args.steps // 2), 1, sample_clocks=False)
        m2 = r2["met"]
        commit_b2 = {"workload": workload_name(log_h, width, 2), "ms_per_step": r2["ms_step"],
                     "value": world * algorithmic_bytes(log_h, width, 2) / (r2["ms_step"] * 1e-3) / 1e9, "unit": UNIT,
                     "stage_ms_per_step": {k: m2[k] / r2["steps"] for k in ("lde_ms", "leaf_hash_ms", "compress_ms")},
                     "root": [int(x) for x in r2["root"]]}
        r2["trace"].free()
        r2["eng"].close()
        trim_pool(cx)
        log(f"commit b=2: {r2['ms_step']:.1f} ms/step")

    # ---- proofs
    seg = keccak = leaf = dag = None
    seg_shapes = [(lh, w, pc, 3) for lh, w, pc in A.segment_shapes(args.segment_scale_log)]
    seg_r1 = None
    if "segment" in legs:
        # the ten-AIR shape round 1 quoted (19.9 ms, 82.8 segments/s), kept so that the two rounds compare on one workload
        r1_shapes = [(lh, w, pc, 3) for lh, w, pc in A.segment_shapes(0, legacy=True)]
        seg_r1, sp = leg_shape(cx, "segment_r1_shape", r1_shapes, 1, args.steps,
                               "round-1 ten-AIR segment shape (openvm_b200/airs.py segment_shapes(legacy=True))", orc)
        sp.close()
        log(f"segment (round-1 shape): {seg_r1['device_ms_per_proof']:.1f} ms/proof")
    # single-engine legs first, the multi-engine legs (segments/s, DAG) last: after K engines have run and gone, single-engine
    # proofs in the same process are 1.5x slower (round-1 shape: 18.9 ms run first, 30 ms run after the 12-engine leg)
    seg_sp = None
    if "segment" in legs:
        seg, seg_sp = leg_shape(cx, "segment", seg_shapes, 1, args.steps,
                                "shape-replay RV32IM segment (per-AIR table of the reference's example, crates/prof/data/"
                                f"metrics.example.json group ecrecover_program, heights x 2^{args.segment_scale_log})",
                                orc, modmul_peak, hbm_peak, cpu_scale_log=-4 if "cpu" in legs else None)
        seg["segments_per_s_single_engine"] = seg.pop("proofs_per_s")
        log(f"segment: {seg['device_ms_per_proof']:.1f} ms/proof on one engine")
    if "keccak" in legs:
        keccak, sp = leg_shape(cx, "keccak", [(17, 3164, 322, 3)], 1, max(2, args.steps // 2),
                               "KeccakVmAir shape (config 4): 2^17 rows x 3164 main columns, 322 EF after-challenge columns",
                               orc, modmul_peak, hbm_peak)
        sp.close()
        log(f"keccak: {keccak['device_ms_per_proof']:.1f} ms/proof")
    if "leaf" in legs:
        lshapes = [(lh, w, None, dg) for lh, w, dg in A.leaf_shapes(args.leaf_scale_log)]
        leaf, sp = leg_shape(cx, "leaf", lshapes, 2, max(2, args.steps // 2),
                             "leaf-aggregation shape (config 5), quotient degree up to 4", orc, modmul_peak, hbm_peak)
        sp.close()
        log(f"leaf: {leaf['device_ms_per_proof']:.1f} ms/proof")
    if seg_sp is not None:
        # engines per GPU: as asked, but no more than two host threads per CPU this rank owns (8 ranks on a 32-CPU host)
        engines = max(1, min(args.proof_engines, 2 * len(cx.cpus)))
        if engines > 1:
            seg.update(leg_segment_throughput(cx, seg_sp, seg_shapes, 1, engines, args.steps))
        else:
            seg["segments_per_s"] = seg["segments_per_s_single_engine"]
        seg["cells_per_s"] = (seg_sp.cells + seg_sp.perm_cells) * seg["segments_per_s"]
        seg_sp.close()
        log(f"segment: {seg['segments_per_s']:.1f} segments/s with {seg.get('engines_per_gpu', 1)} engines per GPU")
    if "dag" in legs:
        agg_shapes = [(lh, w, None, dg) for lh, w, dg in A.leaf_shapes(args.dag_agg_scale_log)]
        dag = leg_dag(cx, args.dag_segments, args.dag_engines, seg_shapes, agg_shapes, orc)
        if dag:
            log(f"dag: {dag['wall_s']:.2f} s for {dag['jobs']} jobs; critical path {dag['critical_path_s']:.2f} s")
    sharded_leg = None
    if "sharded" in legs and world > 1 and world & (world - 1) == 0:
        sharded_leg = leg_sharded(cx, ms_step, b, max(2, args.steps // 2))

    if rank == 0:
        steps = args.steps
        H = 1 << log_h
        rows = H << b
        leaf_ms = met["leaf_hash_ms"] / max(1, met["leaf_hash_launches"])
        leaf_bytes = rows * width * 4 + rows * 32
        lde_bytes = (1 + (1 << b)) * H * width * 4
        lde_ms_step = met["lde_ms"] / steps
        perms = permutations(log_h, width, b)
        roofline = {
            "kernel": "leaf_mixed_kernel (Poseidon2 leaf sponge on the integer-multiply and FP64 pipes side by side; reported "
                      "against HBM as the contract asks, bound by modular products: see int_pipe)",
            "bound": "hbm", "achieved": leaf_bytes / (leaf_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
            "frac": leaf_bytes / (leaf_ms * 1e-3) / 1e9 / hbm_peak,
            # dram__bytes_read.sum + dram__bytes_write.sum of this launch, ncu --set full capture profiles/r1l_ncu_leaf_summary.md
            "traffic": 35.508e9 if (log_h, width, b) == (24, 256, 1) else None,
            "traffic_source": "one ncu --set full capture of this kernel at this shape (profiles/r1l_ncu_leaf_summary.md), "
                              "not re-measured per run",
            "peak_source": peak_src, "ms_per_launch": leaf_ms, "share_of_step": met["leaf_hash_ms"] / steps / ms_step,
        }
        roofline_lde = {
            "kernel": "ntt passes (iNTT + coset NTT)", "bound": "hbm",
            "achieved": lde_bytes / (lde_ms_step * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
            "frac": lde_bytes / (lde_ms_step * 1e-3) / 1e9 / hbm_peak, "ms_per_step": lde_ms_step,
            "launches_per_step": met["lde_launches"] / steps, "share_of_step": lde_ms_step / ms_step,
            # dram__bytes_read.sum + dram__bytes_write.sum of the six passes of one 32-column group (38.8 GB) x 8 groups, from the
            # ncu --set full capture profiles/r2g_lde_shoup.md: three HBM round trips per transform for 51.5 GB algorithmic
            "traffic": 310.4e9 if (log_h, width, b) == (24, 256, 1) else None,
            "traffic_source": "one ncu --set full capture of these kernels at this shape (profiles/r2g_lde_shoup.md), not re-measured per run",
        }
        hash_ms = (met["leaf_hash_ms"] + met["compress_ms"]) / steps
        butterflies = (1 + (1 << b)) * 0.5 * H * log_h * width
        int_pipe = {
            "modmul_per_s_peak": modmul_peak,
            "modmul_peak_source": "b2s_bench_modmul: 8 independent signed-Montgomery squaring chains per thread (the cheapest product "
                                  "form, the S-box's), state in registers, 2048 threads per SM: the one peak every fraction uses",
            "poseidon2_perm_per_s": perms / (hash_ms * 1e-3),
            "poseidon2_perm_per_s_integer_only_microbench": nthreads * 64 / (pms * 1e-3),
            "hash_products_per_s": perms / (hash_ms * 1e-3) * 564,
            "hash_frac_of_modmul_peak": perms / (hash_ms * 1e-3) * 564 / modmul_peak,
            "lde_butterflies_per_s": butterflies / (lde_ms_step * 1e-3),
            "lde_frac_of_modmul_peak": butterflies / (lde_ms_step * 1e-3) / modmul_peak,
            "note": "the hashing runs on the integer AND the FP64 pipe, so its product rate may exceed the integer-only "
                    "Poseidon2 micro-benchmark; fractions are against the modular-product peak",
        }
        cpu, root_checked = None, None
        if "cpu" in legs:
            threads = len(cx.all_cpus)
            os.sched_setaffinity(0, cx.all_cpus)
            r_small, s_small, _ = cpu_commit(orc, small_log, width, b, threads)
            assert list(r_small) == [int(x) for x in small_root], "2^20 root differs from the oracle's"
            root_checked = {"2^%d x %d" % (small_log, width): "oracle (orc_fast_commit)"}
            est_full = s_small * (1 << (log_h - small_log)) * 1.2
            c_log = log_h
            while c_log > small_log and s_small * (1 << (c_log - small_log)) * 1.2 > args.cpu_full_budget_s:
                c_log -= 1
            if c_log == log_h and log_h > small_log:
                r_full, s_full, _ = cpu_commit(orc, log_h, width, b, threads, seed=SEED + rank)
                assert list(r_full) == [int(x) for x in root], "full-size root differs from the oracle's"
                root_checked["2^%d x %d" % (log_h, width)] = "oracle (orc_fast_commit, full size, same seed as the timed steps)"
                cpu_s, sample = s_full, f"the full workload once (2^{log_h} x {width}), {s_full:.1f} s"
            elif c_log > small_log:
                _, cpu_s, _ = cpu_commit(orc, c_log, width, b, threads)
                sample = (f"2^{c_log} x {width} of the same synthetic trace once, {cpu_s:.1f} s (full size estimated at "
                          f"{est_full:.0f} s: over budget)")
            else:
                cpu_s, sample = s_small, f"2^{small_log} x {width} of the same synthetic trace once, {s_small:.2f} s"
            cpu = {"value": algorithmic_bytes(c_log, width, b) / cpu_s / 1e9, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": sample + "; AVX2 Montgomery port of the commit (oracle/fast_commit.cpp), trace generation excluded",
                   "same_config": c_log == log_h}
            log(f"cpu: {cpu['value']:.2f} GB/s on {threads} threads ({sample})")
        launches = met["lde_launches"] + met["leaf_hash_launches"] + met["compress_launches"] + met["other_launches"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic",
            "config": {"workload": workload_name(log_h, width, b), "segments_per_gpu": 1, "cpu_affinity": cx.affinity, "host_wait_policy": cx.sched,
                       "l2": "inputs (16 GiB trace, 32 GiB LDE) are far larger than the 126 MB L2; no flush needed",
                       "fri": {"log_blowup": b}, "legs": sorted(legs)},
            "roofline": roofline, "roofline_lde": roofline_lde, "int_pipe": int_pipe, "cpu_baseline": cpu,
            "root_checked": root_checked, "e2e": e2e, "commit_b2": commit_b2, "segment_proof": seg, "segment_proof_r1_shape": seg_r1, "keccak_proof": keccak,
            "leaf_proof": leaf, "continuation_dag": dag, "sharded_commit": sharded_leg,
            "gpu_launches": int(launches), "clocks": base["clocks"],
            "stage_ms_per_step": {"main_trace_commit_time_ms": met["commit_total_ms"] / steps,
                                  "lde_ms": lde_ms_step, "leaf_hash_ms": met["leaf_hash_ms"] / steps,
                                  "compress_ms": met["compress_ms"] / steps},
            "root": [int(x) for x in root], "reference_example": REFERENCE_EXAMPLE,
            "bench_wall_s": time.perf_counter() - t_start,
        }
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
