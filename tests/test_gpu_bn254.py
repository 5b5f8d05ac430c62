"""GPU parity, root configuration (BabyBearPoseidon2RootConfig): the device Poseidon2-BN254 permutation, the multi-field
leaf sponge and the one-scalar Merkle trees equal the oracle's (oracle/bn254.hpp), through the C ABI."""
import numpy as np
import pytest

import pyref

pytestmark = pytest.mark.gpu
P = 2013265921
rng = np.random.default_rng(23)


def rand(*shape):
    return rng.integers(0, P, size=shape, dtype=np.uint64).astype(np.uint32)


@pytest.fixture
def root_engine(orc):
    from openvm_b200 import B2Engine, FriParameters
    e = B2Engine.new(FriParameters(log_blowup=1))
    e.set_hash(1)
    orc.set_hash(1)
    yield e
    orc.set_hash(0)
    e.close()


def test_device_permutation(engine, orc):
    states = np.frombuffer(rng.bytes(96 * 300), dtype=np.uint32).reshape(-1, 24).copy()
    states[:, 7::8] &= 0x0FFFFFFF          # below the modulus
    states[0] = 0
    states[1] = np.array(pyref.bn_limbs(pyref.BN_R - 1) * 3, dtype=np.uint32)
    states[2] = np.array(pyref.bn_limbs(0) + pyref.bn_limbs(1) + pyref.bn_limbs(2), dtype=np.uint32)
    got = engine.device().poseidon2_bn254_permute(states)
    assert np.array_equal(got, orc.bn254_permute(states))
    # the published known answer, on the device
    want = [0x0bb61d24daca55eebcb1929a82650f328134334da98ea4f847f760054f4a3033,
            0x303b6f7c86d043bfcbcc80214f26a30277a15d3f74ca654992defe7ff8d03570,
            0x1ed25194542b12eef8617361c3ba7c52e660b145994427cc86296242cf766ec8]
    assert [pyref.bn_from_limbs(got[2][8 * k:8 * k + 8]) for k in range(3)] == want


MMCS_CASES = [
    [(1, 1)],
    [(1, 20)],
    [(2, 8)],
    [(8, 3)],
    [(64, 15)],
    [(64, 16)],
    [(64, 17)],
    [(64, 16), (64, 1)],           # sponge block straddles two matrices
    [(16, 5), (4, 2), (4, 7), (1, 3)],
    [(2, 1), (256, 17), (128, 8), (8, 30)],
    [(2048, 33), (256, 70)],
]


@pytest.mark.parametrize("shapes", MMCS_CASES)
def test_mmcs_root_layers_openings(root_engine, orc, shapes):
    dev = root_engine.device()
    mats = [rand(h, w) for h, w in shapes]
    mats[0][0, 0] = P - 1
    dmats = [dev.transport_matrix_to_device(m) for m in mats]
    root, pcs = dev.mmcs_commit(dmats)
    want_root, want_layers = orc.mmcs_commit(mats, want_layers=True)
    assert np.array_equal(root, want_root)
    hmax = max(h for h, _ in shapes)
    off, n, level = 0, hmax, 0
    while n >= 1:
        assert np.array_equal(pcs.layer(level), want_layers[off:off + n]), f"layer {level}"
        off += n
        n //= 2
        level += 1
    for idx in sorted({0, hmax - 1, hmax // 3}):
        rows, proof = pcs.open_batch(idx)
        ok, want_proof = orc.mmcs_open_verify(mats, idx)
        assert ok and np.array_equal(proof, want_proof)
        for m, r in zip(mats, rows):
            lh = m.shape[0].bit_length() - 1
            assert np.array_equal(r, m[idx >> (pcs.log_max_height - lh)])


def test_leaf_is_the_python_sponge(root_engine):
    # independent of the oracle: one row against the pure-Python restatement
    dev = root_engine.device()
    m = rand(1, 37)
    root, _ = dev.mmcs_commit([dev.transport_matrix_to_device(m)])
    assert pyref.bn_from_limbs(root) == pyref.bn_hash_slice([int(v) for v in m[0]])


@pytest.mark.parametrize("shapes", [[(32, 2)], [(1 << 10, 10)], [(1 << 12, 2), (1 << 9, 36), (4, 5), (1, 7)]])
def test_commit_root(root_engine, orc, shapes):
    dev = root_engine.device()
    mats = [rand(h, w) for h, w in shapes]
    root, pcs = dev.commit([dev.transport_matrix_to_device(m) for m in mats])
    assert np.array_equal(root, orc.commit(mats, 1))


def test_streamed_commit_carries_the_sponge_state(root_engine, orc):
    # commit_host hashes column groups as their LDE completes; the row state (three scalars) travels between groups.
    # 2^21 rows make a group 64 columns wide: 150 columns = 64 + 64 + 22. The oracle would need minutes at this height, so
    # the streamed root is compared with the one-shot tree of the same matrix, and a small case with the oracle.
    dev = root_engine.device()
    m = rand(1 << 21, 150)
    root_streamed, pcs = dev.commit_host(m)
    pcs.free()
    root_once, pcs2 = dev.commit([dev.transport_matrix_to_device(m)])
    assert np.array_equal(root_streamed, root_once)
    pcs2.free()
    small = rand(1 << 8, 40)
    root_small, _ = dev.commit_host(small)
    assert np.array_equal(root_small, orc.commit([small], 1))


# ---- whole proofs under the root configuration ----------------------------------------------------------------------------
def _root_engine(orc, b, queries, pow_bits):
    from openvm_b200 import B2Engine, FriParameters
    e = B2Engine.new(FriParameters(log_blowup=b, num_queries=queries, proof_of_work_bits=pow_bits))
    e.set_hash(1)
    orc.set_hash(1)
    return e


def _same_proof(orc, eng, key, inputs, air_objs, b, queries, pow_bits, expect=None):
    from test_gpu_prove import gpu_prove, gpu_verify
    proof = gpu_prove(eng, key, inputs, air_objs)
    res, mine = orc.stark_verify(key, proof, b, queries, pow_bits), gpu_verify(eng, key, proof)
    if expect is None:
        assert res is None and mine is None
    else:
        assert expect in res and expect in mine
    assert proof == orc.stark_prove(key, inputs, b, queries, pow_bits)
    return proof


@pytest.mark.parametrize("n,b,queries,pow_bits", [(32, 1, 4, 3), (8, 2, 3, 0), (1 << 10, 1, 5, 8)])
def test_root_fibonacci_proof_bytes(orc, n, b, queries, pow_bits):
    from openvm_b200 import airs
    from test_oracle_stark import build_key
    eng = _root_engine(orc, b, queries, pow_bits)
    try:
        air = airs.fibonacci_air()
        key = build_key(orc, [air], b)
        trace, pvs = airs.fibonacci_trace(n)
        proof = _same_proof(orc, eng, key, [(0, [trace], pvs)], [air], b, queries, pow_bits)
        # the inner configuration rejects it, on both sides
        orc.set_hash(0)
        assert orc.stark_verify(key, proof, b, queries, pow_bits) is not None
    finally:
        orc.set_hash(0)
        eng.close()


def test_root_multi_air_prep_cached_logup_bytes(orc):
    from openvm_b200 import airs
    from test_oracle_stark import build_key, interaction_traces
    from test_gpu_prove import gpu_prove, gpu_verify
    b, queries, pow_bits = 1, 6, 4
    eng = _root_engine(orc, b, queries, pow_bits)
    try:
        a0 = airs.fibonacci_air()
        a1 = airs.chain_air(7, with_prep=True, cached_width=3, height=64, seed=1)
        a2 = airs.chain_air(5)
        snd, rcv = airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)
        air_objs = [a0, a1, a2, snd, rcv]
        for a in air_objs:
            a.prep_commit = None
        key = build_key(orc, air_objs, b)      # preprocessed commitments under the root hash
        t0, pv0 = airs.fibonacci_trace(256)
        p1, pv1 = airs.chain_trace(a1, 64, seed=5)
        p2, pv2 = airs.chain_trace(a2, 512, seed=6)
        send, recv = interaction_traces(64, 128, 2, seed=4)
        none = np.zeros(0, dtype=np.uint32)
        inputs = [(0, [t0], pv0), (1, p1, pv1), (2, p2, pv2), (3, [send], none), (4, [recv], none)]
        proof = _same_proof(orc, eng, key, inputs, air_objs, b, queries, pow_bits)
        _same_proof(orc, eng, key, inputs[:3], air_objs, b, queries, pow_bits)          # no challenge phase
        send2, recv2 = interaction_traces(64, 128, 2, seed=4, unbalanced=True)
        _same_proof(orc, eng, key, inputs[:3] + [(3, [send2], none), (4, [recv2], none)], air_objs, b, queries, pow_bits,
                    expect="NonZeroCumulativeSum")
        bad = [m.copy() for m in p2]
        bad[0][7, 2] = (int(bad[0][7, 2]) + 1) % P
        _same_proof(orc, eng, key, inputs[:2] + [(2, bad, pv2)] + inputs[3:], air_objs, b, queries, pow_bits,
                    expect="OodEvaluationMismatch")
        # tampered bytes: both verifiers reject
        rng2 = np.random.default_rng(5)
        for pos in rng2.integers(4, len(proof), size=25):
            q = bytearray(proof)
            q[pos] ^= 1 << int(rng2.integers(0, 8))
            assert orc.stark_verify(key, bytes(q), b, queries, pow_bits) is not None
            assert gpu_verify(eng, key, bytes(q)) is not None
        # the inner verifier's hint stream is not defined for a root proof; its counterpart is the halo2 witness
        with pytest.raises(Exception):
            eng.proof_hint_stream(proof)
        for a, b in zip(eng.proof_witness_stream(proof), orc.proof_witness_stream(proof)):
            assert np.array_equal(a, b)
    finally:
        orc.set_hash(0)
        eng.close()
