"""The product's verifier without a device (b2s_verify_host): same verdicts as the oracle's restatement of the in-tree
recursive verifier (extensions/native/recursion/src/stark/mod.rs:108-1173) on oracle-made proofs -- accepted proofs, every
negative case the reference tests (pranked trace -> OodEvaluationMismatch, crates/circuits/poseidon2-air/src/tests.rs:56-74;
unbalanced buses -> NonZeroCumulativeSum), tampered bytes, wrong parameters, both configurations, the key's trace-height
constraints. Runs in the CPU suite: transcript (AVX2 Poseidon2 / BN254), Merkle paths, codec and constraint evaluation of
the shipped library are exercised on every round, not only on the GPU box."""
import numpy as np
import pytest

from openvm_b200 import airs
from test_oracle_stark import build_key, interaction_traces

P = 2013265921
NONE = np.zeros(0, dtype=np.uint32)


def mine(key, proof, b, q, pw, kind=0):
    from openvm_b200 import FriParameters, VerificationError, verify_host
    try:
        verify_host(key, proof, FriParameters(log_blowup=b, num_queries=q, proof_of_work_bits=pw), kind)
        return None
    except VerificationError as ex:
        return str(ex)


def same_verdict(orc, key, proof, b, q, pw, kind=0, expect=None):
    want, got = orc.stark_verify(key, proof, b, q, pw), mine(key, proof, b, q, pw, kind)
    assert (want is None) == (got is None), (want, got)
    if expect is not None:
        assert expect in want and expect in got, (want, got)
    return got


@pytest.fixture(params=[0, 1], ids=["inner", "root"])
def kind(request, orc):
    orc.set_hash(request.param)
    yield request.param
    orc.set_hash(0)


def test_fibonacci_and_negative_cases(orc, kind):
    b, q, pw = 1, 4, 3
    air = airs.fibonacci_air()
    key = build_key(orc, [air], b)
    trace, pvs = airs.fibonacci_trace(32)
    proof = orc.stark_prove(key, [(0, [trace], pvs)], b, q, pw)
    assert same_verdict(orc, key, proof, b, q, pw, kind) is None
    bad = pvs.copy(); bad[2] = (int(bad[2]) + 1) % P
    same_verdict(orc, key, orc.stark_prove(key, [(0, [trace], bad)], b, q, pw), b, q, pw, kind, "OodEvaluationMismatch")
    t2 = trace.copy(); t2[16, 1] = (int(t2[16, 1]) + 1) % P
    same_verdict(orc, key, orc.stark_prove(key, [(0, [t2], pvs)], b, q, pw), b, q, pw, kind, "OodEvaluationMismatch")
    assert same_verdict(orc, key, proof, b, q + 1, pw, kind) is not None          # verifier parameters must match
    assert mine(key, proof, b, q, pw, 1 - kind) is not None                        # and so must the configuration


def test_multi_air_prep_cached_logup(orc, kind):
    b, q, pw = 1, 5, 4
    a0 = airs.fibonacci_air()
    a1 = airs.chain_air(7, with_prep=True, cached_width=3, height=64, seed=1)
    a2 = airs.chain_air(5)
    snd, rcv = airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)
    air_objs = [a0, a1, a2, snd, rcv]
    key = build_key(orc, air_objs, b)
    t0, pv0 = airs.fibonacci_trace(128)
    p1, pv1 = airs.chain_trace(a1, 64, seed=5)
    p2, pv2 = airs.chain_trace(a2, 256, seed=6)
    send, recv = interaction_traces(64, 128, 2, seed=4)
    inputs = [(0, [t0], pv0), (1, p1, pv1), (2, p2, pv2), (3, [send], NONE), (4, [recv], NONE)]
    proof = orc.stark_prove(key, inputs, b, q, pw)
    assert same_verdict(orc, key, proof, b, q, pw, kind) is None
    assert same_verdict(orc, key, orc.stark_prove(key, inputs[:3], b, q, pw), b, q, pw, kind) is None
    send2, recv2 = interaction_traces(64, 128, 2, seed=4, unbalanced=True)
    same_verdict(orc, key, orc.stark_prove(key, inputs[:3] + [(3, [send2], NONE), (4, [recv2], NONE)], b, q, pw), b, q, pw, kind,
                 "NonZeroCumulativeSum")
    rng = np.random.default_rng(3)
    rejected = 0
    for pos in rng.integers(4, len(proof), size=60):
        t = bytearray(proof)
        t[pos] ^= 1 << int(rng.integers(0, 8))
        rejected += same_verdict(orc, key, bytes(t), b, q, pw, kind) is not None
    assert rejected == 60
    assert same_verdict(orc, key, proof[:-1], b, q, pw, kind) is not None


def test_blowup2_and_height_constraints(orc, kind):
    b, q, pw = 2, 3, 2
    a, fib = airs.chain_air(9, degree=3), airs.fibonacci_air()
    parts, pv = airs.chain_trace(a, 64, seed=9)
    t0, pv0 = airs.fibonacci_trace(16)
    inputs = [(0, parts, pv), (1, [t0], pv0)]
    ok = airs.key_blob([a, fib], trace_height_constraints=[([1, 1], 81)])
    tight = airs.key_blob([a, fib], trace_height_constraints=[([1, 1], 80)])
    proof = orc.stark_prove(ok, inputs, b, q, pw)
    assert same_verdict(orc, ok, proof, b, q, pw, kind) is None
    same_verdict(orc, tight, proof, b, q, pw, kind, "trace heights")


def test_bad_arguments():
    from openvm_b200 import B2StarkError, FriParameters, verify_host
    with pytest.raises(B2StarkError):
        verify_host(np.array([1, 2, 3], dtype=np.uint32), b"\x01\x00\x00\x00", FriParameters(log_blowup=1))
    with pytest.raises(B2StarkError):
        verify_host(airs.key_blob([airs.fibonacci_air()]), b"\x01\x00\x00\x00", FriParameters(log_blowup=1), 7)


def test_key_with_malformed_dead_node_is_refused_at_load():
    """A dead DAG node (not reachable from any constraint) with a bad column, operand or kind must fail key loading: the
    verifier evaluates every node, so such a key would prove and then never verify. Same for a non-canonical preprocessed
    commitment word."""
    from openvm_b200 import B2StarkError, FriParameters, verify_host
    fp = FriParameters(log_blowup=1)
    for bad in [(airs.N_MAIN, 999, 0, 0), (airs.N_ADD, 0, 10 ** 6, 0), (99, 0, 0, 0), (airs.N_PUBLIC, 50, 0, 0),
                (airs.N_CONST, P, 0, 0), (airs.N_PERM, 0, 0, 0)]:
        air = airs.fibonacci_air()
        air.dag.nodes.append(bad)
        with pytest.raises(B2StarkError) as ei:
            verify_host(airs.key_blob([air]), b"\x01\x00\x00\x00", fp)
        assert ei.value.status == -1 and "malformed DAG node" in str(ei.value), bad
    a = airs.chain_air(5, with_prep=True, height=16, seed=2)
    a.prep_commit = np.full(8, P, dtype=np.uint32)
    with pytest.raises(B2StarkError) as ei:
        verify_host(airs.key_blob([a]), b"\x01\x00\x00\x00", fp)
    assert "non-canonical preprocessed commitment" in str(ei.value)


def test_decoder_and_key_parser_survive_structured_mutations(orc, kind):
    """Length fields overwritten with extreme values, truncations, insertions, deletions and byte noise in the proof, extreme
    words in the key blob: the library answers with a status (never a crash, a hang or a giant allocation) and accepts a mutant
    only when it is byte-identical to the proof."""
    import struct
    from openvm_b200 import B2StarkError, FriParameters, VerificationError, verify_host
    fib, snd, rcv = airs.fibonacci_air(), airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)
    key = airs.key_blob([fib, snd, rcv], log_up_pow_bits=3)
    send, recv = interaction_traces(16, 32, 2, seed=4)
    t0, pv0 = airs.fibonacci_trace(16)
    proof = orc.stark_prove(key, [(0, [t0], pv0), (1, [send], NONE), (2, [recv], NONE)], 1, 3, 2)
    fp = FriParameters(log_blowup=1, num_queries=3, proof_of_work_bits=2)
    verify_host(key, proof, fp, kind)
    rng = np.random.default_rng(7 + kind)
    extremes = [0, 1, 2, 0xFFFFFFFF, 0x7FFFFFFF, len(proof), len(proof) // 4, 1 << 20]
    for _ in range(600):
        q = bytearray(proof)
        m = int(rng.integers(0, 5))
        if m == 0:
            pos = int(rng.integers(0, len(q) - 4)) & ~3
            q[pos:pos + 4] = struct.pack("<I", int(rng.choice(extremes)))
        elif m == 1:
            q = q[:int(rng.integers(0, len(q)))]
        elif m == 2:
            pos = int(rng.integers(0, len(q)))
            q[pos:pos] = bytes(rng.integers(0, 256, size=int(rng.integers(1, 9)), dtype=np.uint8))
        elif m == 3:
            a = int(rng.integers(0, len(q) - 8))
            del q[a:int(rng.integers(a, min(len(q), a + 64)))]
        else:
            for _ in range(int(rng.integers(1, 6))):
                q[int(rng.integers(0, len(q)))] = int(rng.integers(0, 256))
        try:
            verify_host(key, bytes(q), fp, kind)
            assert bytes(q) == proof
        except VerificationError:
            assert bytes(q) != proof
    for _ in range(300):
        k = key.copy()
        k[int(rng.integers(0, k.size))] = int(rng.choice([0, 1, 2, 0xFFFFFFFF, 0x7FFFFFFF, 1 << 20, k.size]))
        try:
            verify_host(k, proof, fp, kind)
        except (VerificationError, B2StarkError):
            pass
