"""CPU oracle: STARK prove -> codec -> verify, on the reference's own fixture shapes (Fibonacci AIR n = 32,
extensions/native/recursion/src/tests.rs:109) and on the chain AIR; negative tests mirror
crates/circuits/poseidon2-air/src/tests.rs:56-74 (pranked trace -> OodEvaluationMismatch)."""
import numpy as np
import pytest

from openvm_b200 import airs

P = 2013265921


def build_key(orc, air_list, log_blowup):
    for a in air_list:
        if a.prep is not None:
            a.prep_commit = orc.commit([a.prep], log_blowup)
    return airs.key_blob(air_list)


def test_challenger_duplex_order(orc):
    # sample pops from the END of the rate part; an observe after a sample invalidates the remaining outputs
    s = orc.challenger_run(np.arange(1, 9, dtype=np.uint32), 8)
    st = np.zeros((1, 16), dtype=np.uint32)
    st[0, :8] = np.arange(1, 9)
    perm = orc.permute(st)[0]
    assert list(s) == list(perm[:8][::-1])
    # 3 pending inputs: the permutation runs at the first sample
    s2 = orc.challenger_run(np.array([5, 6, 7], dtype=np.uint32), 2)
    st[0] = 0
    st[0, :3] = (5, 6, 7)
    perm2 = orc.permute(st)[0]
    assert list(s2) == [perm2[7], perm2[6]]


@pytest.mark.parametrize("n,b,queries,pow_bits", [(32, 1, 4, 3), (8, 2, 3, 0), (1 << 10, 1, 5, 8)])
def test_fibonacci_prove_verify(orc, n, b, queries, pow_bits):
    air = airs.fibonacci_air()
    key = build_key(orc, [air], b)
    trace, pvs = airs.fibonacci_trace(n)
    proof = orc.stark_prove(key, [(0, [trace], pvs)], b, queries, pow_bits)
    assert orc.stark_verify(key, proof, b, queries, pow_bits) is None
    # deterministic (minimum PoW witness)
    assert proof == orc.stark_prove(key, [(0, [trace], pvs)], b, queries, pow_bits)
    # wrong public value
    bad = pvs.copy(); bad[2] = (int(bad[2]) + 1) % P
    p2 = orc.stark_prove(key, [(0, [trace], bad)], b, queries, pow_bits)
    assert "OodEvaluationMismatch" in orc.stark_verify(key, p2, b, queries, pow_bits)
    # pranked trace cell
    t2 = trace.copy(); t2[n // 2, 1] = (int(t2[n // 2, 1]) + 1) % P
    p3 = orc.stark_prove(key, [(0, [t2], pvs)], b, queries, pow_bits)
    assert "OodEvaluationMismatch" in orc.stark_verify(key, p3, b, queries, pow_bits)
    # verifier parameters must match
    assert orc.stark_verify(key, proof, b, queries + 1, pow_bits) is not None


def test_proof_bytes_tamper(orc):
    air = airs.fibonacci_air()
    key = build_key(orc, [air], 1)
    trace, pvs = airs.fibonacci_trace(64)
    proof = bytearray(orc.stark_prove(key, [(0, [trace], pvs)], 1, 3, 2))
    assert orc.stark_verify(key, bytes(proof), 1, 3, 2) is None
    rng = np.random.default_rng(3)
    rejected = 0
    for pos in rng.integers(4, len(proof), size=40):
        q = bytearray(proof)
        q[pos] ^= 1
        if orc.stark_verify(key, bytes(q), 1, 3, 2) is not None:
            rejected += 1
    assert rejected == 40
    assert orc.stark_verify(key, bytes(proof[:-1]), 1, 3, 2) is not None


def test_multi_air_mixed_heights_prep_cached(orc):
    b = 1
    a0 = airs.fibonacci_air()
    a1 = airs.chain_air(7, with_prep=True, cached_width=3, height=64, seed=1)
    a2 = airs.chain_air(5)
    a3 = airs.chain_air(9, with_prep=True, height=16, seed=2)
    key = build_key(orc, [a0, a1, a2, a3], b)
    t0, pv0 = airs.fibonacci_trace(256)
    p1, pv1 = airs.chain_trace(a1, 64, seed=5)
    p2, pv2 = airs.chain_trace(a2, 512, seed=6)
    p3, pv3 = airs.chain_trace(a3, 16, seed=7)
    inputs = [(0, [t0], pv0), (1, p1, pv1), (2, p2, pv2), (3, p3, pv3)]
    proof = orc.stark_prove(key, inputs, b, 6, 4)
    assert orc.stark_verify(key, proof, b, 6, 4) is None
    # optional AIRs: a subset of the key's AIRs (extensions/native/recursion/src/tests.rs optional-AIR cases)
    proof2 = orc.stark_prove(key, [inputs[1], inputs[3]], b, 6, 4)
    assert orc.stark_verify(key, proof2, b, 6, 4) is None
    # cached partition inconsistent with the common main
    bad = [m.copy() for m in p1]
    bad[0][3, 1] = (int(bad[0][3, 1]) + 5) % P
    proof3 = orc.stark_prove(key, [inputs[0], (1, bad, pv1), inputs[2], inputs[3]], b, 6, 4)
    assert "OodEvaluationMismatch" in orc.stark_verify(key, proof3, b, 6, 4)


def test_blowup2_quotient_degree2(orc):
    b = 2
    a = airs.chain_air(6)
    key = build_key(orc, [a], b)
    parts, pv = airs.chain_trace(a, 128, seed=9)
    proof = orc.stark_prove(key, [(0, parts, pv)], b, 4, 5)
    assert orc.stark_verify(key, proof, b, 4, 5) is None


def interaction_traces(n_send, n_recv, n_fields, seed=0, unbalanced=False):
    """sender rows (count, fields) and a receiver holding the same multiset in a different order / grouping."""
    rng = np.random.default_rng(seed)
    send = np.zeros((n_send, 1 + n_fields), dtype=np.uint32)
    send[:, 0] = rng.integers(0, 4, size=n_send)
    send[:, 1:] = rng.integers(0, 50, size=(n_send, n_fields))
    # receiver: merge equal messages, pad with zero-count rows
    agg = {}
    for row in send:
        key = tuple(int(x) for x in row[1:])
        agg[key] = agg.get(key, 0) + int(row[0])
    items = [(c, k) for k, c in agg.items() if c]
    assert len(items) <= n_recv
    recv = np.zeros((n_recv, 1 + n_fields), dtype=np.uint32)
    for i, (c, k) in enumerate(reversed(items)):
        recv[i, 0] = c
        recv[i, 1:] = k
    if unbalanced:
        recv[0, 0] += 1
    return send, recv


def test_logup_sender_receiver(orc):
    b = 1
    snd = airs.dummy_interaction_air(2, True)
    rcv = airs.dummy_interaction_air(2, False)
    fib = airs.fibonacci_air()
    key = airs.key_blob([fib, snd, rcv], log_up_pow_bits=5)
    send, recv = interaction_traces(64, 128, 2, seed=4)
    t0, pv0 = airs.fibonacci_trace(32)
    none = np.zeros(0, dtype=np.uint32)
    inputs = [(0, [t0], pv0), (1, [send], none), (2, [recv], none)]
    proof = orc.stark_prove(key, inputs, b, 5, 4)
    assert orc.stark_verify(key, proof, b, 5, 4) is None
    # without the interacting AIRs there is no challenge phase at all
    p0 = orc.stark_prove(key, inputs[:1], b, 5, 4)
    assert orc.stark_verify(key, p0, b, 5, 4) is None
    # unbalanced buses: the proof is well-formed but the cumulative sums do not cancel
    send2, recv2 = interaction_traces(64, 128, 2, seed=4, unbalanced=True)
    p2 = orc.stark_prove(key, [inputs[0], (1, [send2], none), (2, [recv2], none)], b, 5, 4)
    assert "NonZeroCumulativeSum" in orc.stark_verify(key, p2, b, 5, 4)


def test_logup_two_interactions_per_chunk_with_constraints(orc):
    b = 1
    a0 = airs.lookup_chain_air(6)
    a1 = airs.lookup_chain_receiver()
    key = airs.key_blob([a0, a1], log_up_pow_bits=3)
    parts, pv = airs.chain_trace(a0, 64, seed=3)
    x = parts[-1]
    recv = np.zeros((64, 4), dtype=np.uint32)
    recv[:, 0] = 1
    recv[:, 1:4] = x[::-1, 0:3]
    none = np.zeros(0, dtype=np.uint32)
    proof = orc.stark_prove(key, [(0, parts, pv), (1, [recv], none)], b, 4, 2)
    assert orc.stark_verify(key, proof, b, 4, 2) is None
    bad = recv.copy(); bad[5, 2] = (int(bad[5, 2]) + 1) % P
    p2 = orc.stark_prove(key, [(0, parts, pv), (1, [bad], none)], b, 4, 2)
    assert "NonZeroCumulativeSum" in orc.stark_verify(key, p2, b, 4, 2)


@pytest.mark.parametrize("n,b", [(64, 1), (256, 2)])
def test_logup_prep_cached_rotated_fields(orc, n, b):
    a0 = airs.lookup_chain_air_parts(6, n, seed=2)
    a1 = airs.multi_receiver_air([2, 3], 5)
    key = build_key(orc, [a0, a1], b)
    parts, pv = airs.chain_trace(a0, n, seed=8)
    recv = airs.lookup_chain_parts_receiver_trace(a0, parts)
    none = np.zeros(0, dtype=np.uint32)
    proof = orc.stark_prove(key, [(0, parts, pv), (1, [recv], none)], b, 4, 3)
    assert orc.stark_verify(key, proof, b, 4, 3) is None
    bad = recv.copy(); bad[3, 6] = (int(bad[3, 6]) + 1) % P
    p2 = orc.stark_prove(key, [(0, parts, pv), (1, [bad], none)], b, 4, 3)
    assert "NonZeroCumulativeSum" in orc.stark_verify(key, p2, b, 4, 3)


# ---- quotient degree 4 (constraint degree 5; three interactions per LogUp chunk) -- the leaf-aggregation shape -----------
def multi_bus_traces(field_lens, n, seed=0):
    """rows [m_0, fields_0.., m_1, ...] for multi_receiver_air as a sender, and the same rows in reverse order for the
    receiver: every bus balances."""
    rng = np.random.default_rng(seed)
    w = sum(field_lens) + len(field_lens)
    t = rng.integers(0, 60, size=(n, w)).astype(np.uint32)
    col = 0
    for ln in field_lens:
        t[:, col] = rng.integers(0, 3, size=n)
        col += 1 + ln
    return t, t[::-1].copy()


@pytest.mark.parametrize("b", [2, 3])
def test_quotient_degree4_chain(orc, b):
    a = airs.chain_air(8, degree=5)
    assert a.log_qd == 2
    key = build_key(orc, [a], b)
    parts, pv = airs.chain_trace(a, 64, seed=11)
    proof = orc.stark_prove(key, [(0, parts, pv)], b, 4, 3)
    assert orc.stark_verify(key, proof, b, 4, 3) is None
    bad = [m.copy() for m in parts]
    bad[-1][9, 7] = (int(bad[-1][9, 7]) + 1) % P
    p2 = orc.stark_prove(key, [(0, bad, pv)], b, 4, 3)
    assert "OodEvaluationMismatch" in orc.stark_verify(key, p2, b, 4, 3)


def test_quotient_degree_exceeds_blowup_is_refused(orc):
    a = airs.chain_air(8, degree=5)
    key = build_key(orc, [a], 1)
    parts, pv = airs.chain_trace(a, 64, seed=11)
    with pytest.raises(Exception):
        orc.stark_prove(key, [(0, parts, pv)], 1, 4, 3)


def test_logup_three_per_chunk_quotient_degree4(orc):
    b = 2
    lens = [2, 1, 3, 2]
    snd = airs.multi_receiver_air(lens, 7, max_chunk=3, is_send=True, log_qd=2)
    rcv = airs.multi_receiver_air(lens, 7, max_chunk=3, is_send=False, log_qd=2)
    assert [len(c) for c in snd.chunks] == [3, 1]
    key = airs.key_blob([snd, rcv], log_up_pow_bits=4)
    s, r = multi_bus_traces(lens, 128, seed=5)
    none = np.zeros(0, dtype=np.uint32)
    proof = orc.stark_prove(key, [(0, [s], none), (1, [r], none)], b, 4, 3)
    assert orc.stark_verify(key, proof, b, 4, 3) is None
    r2 = r.copy(); r2[17, 0] += 1          # one more copy of a message than was sent
    p2 = orc.stark_prove(key, [(0, [s], none), (1, [r2], none)], b, 4, 3)
    assert "NonZeroCumulativeSum" in orc.stark_verify(key, p2, b, 4, 3)


def test_trace_height_constraints(orc):
    """LinearConstraint list of the verifying key (extensions/native/recursion/src/types.rs:70): the verifier rejects heights
    with sum coeff_i * height_i >= threshold (stark/mod.rs:836-901), the proving side refuses them (vm.rs:627-652)."""
    fib, ch = airs.fibonacci_air(), airs.chain_air(5)
    t0, pv0 = airs.fibonacci_trace(32)
    p1, pv1 = airs.chain_trace(ch, 64, seed=6)
    inputs = [(0, [t0], pv0), (1, p1, pv1)]
    free = airs.key_blob([fib, ch])
    proof = orc.stark_prove(free, inputs, 1, 3, 2)
    # 1 * 32 + 2 * 64 = 160 < 161: satisfied, and the same proof (the constraints are not part of the transcript)
    ok = airs.key_blob([fib, ch], trace_height_constraints=[([1, 2], 161), ([0, 1], P)])
    assert orc.stark_prove(ok, inputs, 1, 3, 2) == proof
    assert orc.stark_verify(ok, proof, 1, 3, 2) is None
    # 160 >= 160: violated; an absent AIR contributes nothing
    tight = airs.key_blob([fib, ch], trace_height_constraints=[([0, 0], 1), ([1, 2], 160)])
    assert "trace heights" in orc.stark_verify(tight, proof, 1, 3, 2)
    with pytest.raises(RuntimeError, match="linear constraint"):
        orc.stark_prove(tight, inputs, 1, 3, 2)
    only_fib = orc.stark_prove(tight, inputs[:1], 1, 3, 2)
    assert orc.stark_verify(tight, only_fib, 1, 3, 2) is None
    # large coefficients do not wrap: 2^31 * 64 is far above any threshold
    big = airs.key_blob([fib, ch], trace_height_constraints=[([0, 1 << 31], P)])
    assert "trace heights" in orc.stark_verify(big, proof, 1, 3, 2)
    with pytest.raises(RuntimeError):
        airs_bad = airs.key_blob([fib, ch], trace_height_constraints=[([1, 1], P + 1)])
        orc.stark_prove(airs_bad, inputs, 1, 3, 2)
