"""GPU parity, full proofs: bytes from b2s_prove equal the CPU oracle prover's bytes (same minimum PoW witnesses)
and the oracle's verifier (transliterated from the in-tree recursive verifier) accepts them."""
import numpy as np
import pytest

from openvm_b200 import airs
from test_oracle_stark import build_key

pytestmark = pytest.mark.gpu
P = 2013265921


def make_engine(b, queries, pow_bits):
    from openvm_b200 import B2Engine, FriParameters
    return B2Engine.new(FriParameters(log_blowup=b, num_queries=queries, proof_of_work_bits=pow_bits))


def gpu_verify(eng, key, proof):
    """product-side verifier (b2s_verify): None or the VerificationError text"""
    from openvm_b200 import VerificationError
    pk = eng.device().transport_pk_to_device(key)
    try:
        eng.verify(pk, proof)
        return None
    except VerificationError as ex:
        return str(ex)
    finally:
        pk.free()


def gpu_prove(eng, key, inputs, air_objs):
    from openvm_b200 import AirProvingContext
    dev = eng.device()
    pk = dev.transport_pk_to_device(key)
    ctx, keep = [], []
    for air_id, mats, pvs in inputs:
        a = air_objs[air_id]
        cached, traces = [], []
        for m in mats[:len(a.cached_widths)]:
            dm = dev.transport_matrix_to_device(m)
            _, pcs = dev.commit([dm])
            cached.append(pcs)
            traces.append(dm)
        common = dev.transport_matrix_to_device(mats[-1])
        keep.append(common)
        ctx.append(AirProvingContext(air_id, common, pvs, tuple(cached), tuple(traces) if a.interactions else ()))
    proof = eng.prove(pk, ctx)
    pk.free()
    return proof


@pytest.mark.parametrize("n,b,queries,pow_bits", [(32, 1, 4, 3), (8, 2, 3, 0), (1 << 10, 1, 5, 8), (1 << 15, 1, 100, 16),
                                                  (1 << 16, 1, 4, 3)])   # LDE height 2^17: the tall-height kernels
def test_fibonacci_proof_bytes(orc, n, b, queries, pow_bits):
    air = airs.fibonacci_air()
    key = build_key(orc, [air], b)
    trace, pvs = airs.fibonacci_trace(n)
    eng = make_engine(b, queries, pow_bits)
    try:
        proof = gpu_prove(eng, key, [(0, [trace], pvs)], [air])
    finally:
        eng.close()
    assert orc.stark_verify(key, proof, b, queries, pow_bits) is None
    if n <= 1 << 10:
        assert proof == orc.stark_prove(key, [(0, [trace], pvs)], b, queries, pow_bits)


def test_multi_air_proof_bytes(orc):
    b, queries, pow_bits = 1, 6, 4
    a0 = airs.fibonacci_air()
    a1 = airs.chain_air(7, with_prep=True, cached_width=3, height=64, seed=1)
    a2 = airs.chain_air(5)
    a3 = airs.chain_air(9, with_prep=True, height=16, seed=2)
    air_objs = [a0, a1, a2, a3]
    key = build_key(orc, air_objs, b)
    t0, pv0 = airs.fibonacci_trace(256)
    p1, pv1 = airs.chain_trace(a1, 64, seed=5)
    p2, pv2 = airs.chain_trace(a2, 512, seed=6)
    p3, pv3 = airs.chain_trace(a3, 16, seed=7)
    inputs = [(0, [t0], pv0), (1, p1, pv1), (2, p2, pv2), (3, p3, pv3)]
    eng = make_engine(b, queries, pow_bits)
    try:
        proof = gpu_prove(eng, key, inputs, air_objs)
        assert orc.stark_verify(key, proof, b, queries, pow_bits) is None
        assert proof == orc.stark_prove(key, inputs, b, queries, pow_bits)
        sub = [inputs[1], inputs[3]]
        proof2 = gpu_prove(eng, key, sub, air_objs)
        assert proof2 == orc.stark_prove(key, sub, b, queries, pow_bits)
        # pranked trace: the device prover still produces a proof, the verifier rejects it at the OOD check
        bad = [m.copy() for m in p2]
        bad[0][7, 2] = (int(bad[0][7, 2]) + 1) % P
        proof3 = gpu_prove(eng, key, [inputs[0], inputs[1], (2, bad, pv2), inputs[3]], air_objs)
        assert "OodEvaluationMismatch" in orc.stark_verify(key, proof3, b, queries, pow_bits)
        # the library's own verifier agrees with the restated reference verifier
        assert gpu_verify(eng, key, proof) is None and gpu_verify(eng, key, proof2) is None
        assert "OodEvaluationMismatch" in gpu_verify(eng, key, proof3)
    finally:
        eng.close()


def test_blowup2_wide_chain(orc):
    b, queries, pow_bits = 2, 4, 5
    a = airs.chain_air(40)
    key = build_key(orc, [a], b)
    parts, pv = airs.chain_trace(a, 1 << 11, seed=9)
    eng = make_engine(b, queries, pow_bits)
    try:
        proof = gpu_prove(eng, key, [(0, parts, pv)], [a])
    finally:
        eng.close()
    assert orc.stark_verify(key, proof, b, queries, pow_bits) is None
    assert proof == orc.stark_prove(key, [(0, parts, pv)], b, queries, pow_bits)


# ---- FRI-LogUp challenge phase ------------------------------------------------------------------------------------------
NONE = np.zeros(0, dtype=np.uint32)


def check_same(orc, eng, key, inputs, air_objs, b, queries, pow_bits, expect=None):
    proof = gpu_prove(eng, key, inputs, air_objs)
    res = orc.stark_verify(key, proof, b, queries, pow_bits)
    mine = gpu_verify(eng, key, proof)
    if expect is None:
        assert res is None and mine is None
    else:
        assert expect in res and expect in mine
    assert proof == orc.stark_prove(key, inputs, b, queries, pow_bits)
    # what the next aggregation layer would be fed: the hint stream of the device-made proof (hints.rs:272-316)
    assert [list(map(int, v)) for v in eng.proof_hint_stream(proof)] == orc.proof_hint_stream(proof)
    return proof


def test_logup_sender_receiver_bytes(orc):
    from test_oracle_stark import interaction_traces
    b, queries, pow_bits = 1, 5, 4
    air_objs = [airs.fibonacci_air(), airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)]
    key = airs.key_blob(air_objs, log_up_pow_bits=5)
    send, recv = interaction_traces(64, 128, 2, seed=4)
    t0, pv0 = airs.fibonacci_trace(32)
    inputs = [(0, [t0], pv0), (1, [send], NONE), (2, [recv], NONE)]
    eng = make_engine(b, queries, pow_bits)
    try:
        check_same(orc, eng, key, inputs, air_objs, b, queries, pow_bits)
        check_same(orc, eng, key, inputs[:1], air_objs, b, queries, pow_bits)   # no interacting AIR -> no phase
        send2, recv2 = interaction_traces(64, 128, 2, seed=4, unbalanced=True)
        check_same(orc, eng, key, [inputs[0], (1, [send2], NONE), (2, [recv2], NONE)], air_objs, b, queries, pow_bits,
                   expect="NonZeroCumulativeSum")
    finally:
        eng.close()


def test_logup_two_interactions_per_chunk_bytes(orc):
    b, queries, pow_bits = 1, 4, 2
    air_objs = [airs.lookup_chain_air(6), airs.lookup_chain_receiver()]
    key = airs.key_blob(air_objs, log_up_pow_bits=3)
    parts, pv = airs.chain_trace(air_objs[0], 64, seed=3)
    recv = np.zeros((64, 4), dtype=np.uint32)
    recv[:, 0] = 1
    recv[:, 1:4] = parts[-1][::-1, 0:3]
    eng = make_engine(b, queries, pow_bits)
    try:
        check_same(orc, eng, key, [(0, parts, pv), (1, [recv], NONE)], air_objs, b, queries, pow_bits)
    finally:
        eng.close()


@pytest.mark.parametrize("n,b", [(64, 1), (1 << 13, 2)])
def test_logup_prep_cached_rotated_fields_bytes(orc, n, b):
    queries, pow_bits = 4, 3
    a0 = airs.lookup_chain_air_parts(6, n, seed=2)
    a1 = airs.multi_receiver_air([2, 3], 5)
    air_objs = [a0, a1]
    key = build_key(orc, air_objs, b)
    parts, pv = airs.chain_trace(a0, n, seed=8)
    recv = airs.lookup_chain_parts_receiver_trace(a0, parts)
    eng = make_engine(b, queries, pow_bits)
    try:
        check_same(orc, eng, key, [(0, parts, pv), (1, [recv], NONE)], air_objs, b, queries, pow_bits)
    finally:
        eng.close()


# ---- every quotient code path against the ORACLE's bytes -----------------------------------------------------------------
# (a) generated code split into several kernels (the accumulator plane carries the fold from kernel to kernel, qjit.cu), as
#     bench.py's segment / Keccak shapes run it; (b) the interpreter (B2S_QUOTIENT_JIT=0). Both environment variables are read
#     when the key is loaded. The metrics say which kernels actually ran.
def _prove_lookup_parts(orc, monkeypatch, env, n, b):
    queries, pow_bits = 4, 3
    a0 = airs.lookup_chain_air_parts(6, n, seed=2)     # LogUp + preprocessed + cached partitions + rotated fields
    a1 = airs.multi_receiver_air([2, 3], 5)
    a2 = airs.chain_air(40)                            # widest byte-parity AIR, base field only
    air_objs = [a0, a1, a2]
    key = build_key(orc, air_objs, b)
    parts, pv = airs.chain_trace(a0, n, seed=8)
    recv = airs.lookup_chain_parts_receiver_trace(a0, parts)
    p2, pv2 = airs.chain_trace(a2, 2 * n, seed=9)
    inputs = [(0, parts, pv), (1, [recv], NONE), (2, p2, pv2)]
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    eng = make_engine(b, queries, pow_bits)
    try:
        eng.reset_metrics()
        proof = gpu_prove(eng, key, inputs, air_objs)
        m = eng.metrics()
        assert gpu_verify(eng, key, proof) is None
    finally:
        eng.close()
    assert orc.stark_verify(key, proof, b, queries, pow_bits) is None
    assert proof == orc.stark_prove(key, inputs, b, queries, pow_bits)
    return m


@pytest.mark.parametrize("n,b", [(64, 1), (1 << 12, 2)])
def test_split_generated_quotient_kernels_bytes(orc, monkeypatch, n, b):
    m = _prove_lookup_parts(orc, monkeypatch, {"B2S_QUOTIENT_JIT_LINES": "150"}, n, b)
    assert m["quotient_interpreter_launches"] == 0
    assert m["quotient_jit_launches"] >= 12, m      # three AIRs, each split into several kernels
    m1 = _prove_lookup_parts(orc, monkeypatch, {"B2S_QUOTIENT_JIT_LINES": "1000000"}, n, b)
    assert m1["quotient_jit_launches"] == 3, m1      # one kernel per AIR when nothing forces a split


@pytest.mark.parametrize("n,b", [(64, 1), (1 << 12, 2)])
def test_interpreted_quotient_bytes(orc, monkeypatch, n, b):
    m = _prove_lookup_parts(orc, monkeypatch, {"B2S_QUOTIENT_JIT": "0"}, n, b)
    assert m["quotient_jit_launches"] == 0 and m["quotient_interpreter_launches"] == 3, m


def test_verify_agrees_with_oracle_on_tampered_proofs(orc):
    b, queries, pow_bits = 1, 3, 2
    air_objs = [airs.lookup_chain_air(6), airs.lookup_chain_receiver()]
    key = airs.key_blob(air_objs, log_up_pow_bits=3)
    parts, pv = airs.chain_trace(air_objs[0], 64, seed=3)
    recv = np.zeros((64, 4), dtype=np.uint32)
    recv[:, 0] = 1
    recv[:, 1:4] = parts[-1][::-1, 0:3]
    eng = make_engine(b, queries, pow_bits)
    try:
        proof = gpu_prove(eng, key, [(0, parts, pv), (1, [recv], NONE)], air_objs)
        pk = eng.device().transport_pk_to_device(key)
        from openvm_b200 import VerificationError
        eng.verify(pk, proof)
        rng = np.random.default_rng(11)
        for pos in rng.integers(4, len(proof), size=60):
            q = bytearray(proof)
            q[pos] ^= 1 << int(rng.integers(0, 8))
            want = orc.stark_verify(key, bytes(q), b, queries, pow_bits)
            try:
                eng.verify(pk, bytes(q))
                got = None
            except VerificationError as ex:
                got = str(ex)
            assert (want is None) == (got is None), (pos, want, got)
            if want is not None:
                kind = want.split(":")[0]
                kind = "InvalidProofShape" if kind == "proof" else kind    # the oracle's decoder errors
                assert got.startswith(kind), (pos, want, got)
        with pytest.raises(VerificationError):
            eng.verify(pk, proof[:-1])
        # other FRI parameters: reject
        eng2 = make_engine(b, queries + 1, pow_bits)
        try:
            pk2 = eng2.device().transport_pk_to_device(key)
            with pytest.raises(VerificationError):
                eng2.verify(pk2, proof)
            pk2.free()
        finally:
            eng2.close()
        pk.free()
    finally:
        eng.close()


def test_pow_witness_override(orc):
    """replay aid: any valid witness is accepted; the default is the minimum one"""
    b, queries, pow_bits = 1, 3, 6
    air = airs.fibonacci_air()
    key = build_key(orc, [air], b)
    trace, pvs = airs.fibonacci_trace(64)
    eng = make_engine(b, queries, pow_bits)
    try:
        proof = gpu_prove(eng, key, [(0, [trace], pvs)], [air])
        assert gpu_verify(eng, key, proof) is None
        # the default is the minimum witness: every smaller candidate is rejected, some larger one is accepted too
        from openvm_b200 import B2StarkError
        found = None
        for cand in range(0, 4096):
            eng.set_pow_witnesses(None, cand)
            try:
                p2 = gpu_prove(eng, key, [(0, [trace], pvs)], [air])
            except B2StarkError:
                continue
            if p2 != proof:
                found = cand
                assert gpu_verify(eng, key, p2) is None
                assert orc.stark_verify(key, p2, b, queries, pow_bits) is None
                break
        assert found is not None
        eng.set_pow_witnesses(None, None)
        assert gpu_prove(eng, key, [(0, [trace], pvs)], [air]) == proof
    finally:
        eng.close()


# ---- quotient degree 4 at log_blowup 2 / 3 (the leaf-aggregation configuration) ------------------------
@pytest.mark.parametrize("b,n", [(2, 64), (2, 1 << 12), (3, 256)])
def test_quotient_degree4_chain_bytes(orc, b, n):
    a = airs.chain_air(8, degree=5)
    key = build_key(orc, [a], b)
    parts, pv = airs.chain_trace(a, n, seed=11)
    eng = make_engine(b, 4, 3)
    try:
        check_same(orc, eng, key, [(0, parts, pv)], [a], b, 4, 3)
        bad = [m.copy() for m in parts]
        bad[-1][9, 7] = (int(bad[-1][9, 7]) + 1) % P
        check_same(orc, eng, key, [(0, bad, pv)], [a], b, 4, 3, expect="OodEvaluationMismatch")
    finally:
        eng.close()


def test_quotient_degree_exceeds_blowup_is_refused(orc):
    from openvm_b200 import B2StarkError
    a = airs.chain_air(8, degree=5)
    key = build_key(orc, [a], 1)
    eng = make_engine(1, 4, 3)
    try:
        with pytest.raises(B2StarkError):
            eng.device().transport_pk_to_device(key)
    finally:
        eng.close()


def test_mixed_quotient_degrees_with_logup_bytes(orc):
    """one proof over AIRs of quotient degree 1, 2 and 4, two of them with three interactions per LogUp chunk"""
    from test_oracle_stark import multi_bus_traces
    b, queries, pow_bits = 2, 5, 4
    lens = [2, 1, 3, 2]
    air_objs = [airs.fibonacci_air(), airs.chain_air(7), airs.chain_air(9, degree=5),
                airs.multi_receiver_air(lens, 7, max_chunk=3, is_send=True, log_qd=2),
                airs.multi_receiver_air(lens, 7, max_chunk=3, is_send=False, log_qd=2)]
    key = airs.key_blob(air_objs, log_up_pow_bits=4)
    t0, pv0 = airs.fibonacci_trace(128)
    p1, pv1 = airs.chain_trace(air_objs[1], 1 << 9, seed=1)
    p2, pv2 = airs.chain_trace(air_objs[2], 1 << 7, seed=2)
    s, r = multi_bus_traces(lens, 1 << 8, seed=5)
    inputs = [(0, [t0], pv0), (1, p1, pv1), (2, p2, pv2), (3, [s], NONE), (4, [r], NONE)]
    eng = make_engine(b, queries, pow_bits)
    try:
        check_same(orc, eng, key, inputs, air_objs, b, queries, pow_bits)
        r2 = r.copy(); r2[17, 0] += 1
        check_same(orc, eng, key, inputs[:4] + [(4, [r2], NONE)], air_objs, b, queries, pow_bits,
                   expect="NonZeroCumulativeSum")
    finally:
        eng.close()


def test_trace_height_constraints_bytes_and_rejection(orc):
    """The key's LinearConstraint list (types.rs:70): b2s_prove refuses heights that break it (vm.rs:627-652), b2s_verify
    rejects proofs that claim such heights (stark/mod.rs:836-901); same verdicts as the oracle."""
    from openvm_b200 import B2StarkError
    b, queries, pow_bits = 1, 3, 2
    fib, ch = airs.fibonacci_air(), airs.chain_air(5)
    air_objs = [fib, ch]
    t0, pv0 = airs.fibonacci_trace(32)
    p1, pv1 = airs.chain_trace(ch, 64, seed=6)
    inputs = [(0, [t0], pv0), (1, p1, pv1)]
    ok = airs.key_blob(air_objs, trace_height_constraints=[([1, 2], 161), ([0, 1], P)])
    tight = airs.key_blob(air_objs, trace_height_constraints=[([0, 0], 1), ([1, 2], 160)])
    eng = make_engine(b, queries, pow_bits)
    try:
        proof = check_same(orc, eng, ok, inputs, air_objs, b, queries, pow_bits)
        assert proof == gpu_prove(eng, airs.key_blob(air_objs), inputs, air_objs)      # not part of the transcript
        assert "trace heights" in gpu_verify(eng, tight, proof) and "trace heights" in orc.stark_verify(tight, proof, b, queries, pow_bits)
        with pytest.raises(B2StarkError, match="linear constraint 1"):
            gpu_prove(eng, tight, inputs, air_objs)
        only_fib = gpu_prove(eng, tight, inputs[:1], air_objs)
        assert gpu_verify(eng, tight, only_fib) is None
        with pytest.raises(B2StarkError):
            eng.device().transport_pk_to_device(airs.key_blob(air_objs, trace_height_constraints=[([1, 1], P + 1)]))
    finally:
        eng.close()


def test_prover_pool_concurrent_engines_give_the_oracle_bytes(orc):
    """Several engines on one GPU, one host thread each (openvm_b200.scheduler.ProverPool): proofs made concurrently are the
    proofs made alone, and a dependent job receives its parent's proof."""
    from openvm_b200 import AirProvingContext, B2Engine, FriParameters, scheduler
    b, queries, pow_bits = 1, 4, 3
    air = airs.fibonacci_air()
    key = build_key(orc, [air], b)
    sizes = [32, 64, 128, 256, 512, 1024]
    want = {n: orc.stark_prove(key, [(0, [airs.fibonacci_trace(n)[0]], airs.fibonacci_trace(n)[1])], b, queries, pow_bits) for n in sizes}

    class Worker:
        def __init__(self, k):
            self.eng = B2Engine.new(FriParameters(log_blowup=b, num_queries=queries, proof_of_work_bits=pow_bits))
            self.pk = self.eng.device().transport_pk_to_device(key)

        def prove(self, n):
            trace, pvs = airs.fibonacci_trace(n)
            return self.eng.prove(self.pk, [AirProvingContext(0, self.eng.device().transport_matrix_to_device(trace), pvs)])

        def check(self, proof):
            self.eng.verify(self.pk, proof)
            return len(proof)

    workers = []

    def make(k):
        w = Worker(k)
        workers.append(w)
        return w

    pool = scheduler.ProverPool(make, engines=3)
    try:
        jobs = {n: pool.submit(lambda w, n=n: w.prove(n)) for n in sizes for _ in range(1)}
        checks = {n: pool.submit(lambda w, proof: w.check(proof), deps=[j]) for n, j in jobs.items()}
        for n in sizes:
            assert pool.result(jobs[n], timeout=120) == want[n]
            assert pool.result(checks[n], timeout=120) == len(want[n])
    finally:
        pool.close()
        for w in workers:
            w.eng.close()


# ---- zero-copy hand-off from a device tracegen ------------------------------------------------------------
def test_prove_from_adopted_device_traces_on_the_callers_stream(orc):
    """The reference's GPU flow: the trace is written by the caller's kernels into the caller's device memory and handed over
    as DeviceMatrix{buffer, height, width} (crates/vm/src/system/cuda/program.rs:60,119-125; crates/vm/src/arch/vm.rs:588-599).
    Here torch ops on a torch stream play the tracegen; the engine shares that stream (b2s_engine_set_stream), adopts the
    buffers (b2s_matrix_adopt_device) and proves -- no copy and no host synchronisation between generation and proof. The
    proof equals the oracle's proof of the same trace byte for byte."""
    import torch
    from openvm_b200 import AirProvingContext
    b, queries, pow_bits = 1, 4, 3
    air_objs = [airs.segment_air(8, perm_cols=3), airs.segment_air(12, perm_cols=4)]
    heights = [1 << 9, 1 << 6]
    key = airs.key_blob(air_objs, log_up_pow_bits=3)
    eng = make_engine(b, queries, pow_bits)
    try:
        dev = eng.device()
        pk = dev.transport_pk_to_device(key)
        s = torch.cuda.Stream()
        eng.set_stream(s.cuda_stream)
        assert eng.stream == s.cuda_stream
        with torch.cuda.stream(s):
            gen = [airs.chain_trace_device(a, h, seed=31 + i, device="cuda") for i, (a, h) in enumerate(zip(air_objs, heights))]
            ctx = []
            for i, ((cols, pv), a, h) in enumerate(zip(gen, air_objs, heights)):
                m = dev.adopt_device_matrix(cols.data_ptr(), h, a.width_common)
                assert m.device_ptr == cols.data_ptr()
                ctx.append(AirProvingContext(i, m, pv))
            proof = eng.prove(pk, ctx)                       # enqueued behind the generation on the same stream
        inputs = []
        for i, (a, h) in enumerate(zip(air_objs, heights)):
            parts, pv = airs.chain_trace(a, h, seed=31 + i, hashed=True)
            assert np.array_equal(pv, gen[i][1])
            assert np.array_equal(ctx[i].common_main.to_host(), parts[-1])   # what the device generator wrote
            inputs.append((i, parts, pv))
        assert proof == orc.stark_prove(key, inputs, b, queries, pow_bits)
        assert orc.stark_verify(key, proof, b, queries, pow_bits) is None
        pk.free()
    finally:
        eng.close()


@pytest.mark.parametrize("heights", [(2, 4), (8, 2), (16, 1 << 10)])
def test_tiny_trace_heights_bytes(orc, heights):
    """heights of 2, 4, 8 rows next to taller AIRs (the reference's example segment has AIRs of 2, 8 and 16 rows)"""
    b, queries, pow_bits = 1, 4, 2
    air_objs = [airs.segment_air(5, perm_cols=3), airs.segment_air(9, perm_cols=2)]
    key = airs.key_blob(air_objs, log_up_pow_bits=2)
    inputs = []
    for i, (a, h) in enumerate(zip(air_objs, heights)):
        parts, pv = airs.chain_trace(a, h, seed=3 + i)
        inputs.append((i, parts, pv))
    eng = make_engine(b, queries, pow_bits)
    try:
        check_same(orc, eng, key, inputs, air_objs, b, queries, pow_bits)
    finally:
        eng.close()
