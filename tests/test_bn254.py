"""Root configuration hashing (BabyBearPoseidon2RootConfig, crates/continuations/src/lib.rs:17): Poseidon2 over the BN254
scalar field, the multi-field sponge, compression and challenger. CPU side: the published known answer pins the
permutation; the oracle, the pure-Python restatement and the product's host code must agree on everything else."""
import ctypes
import os
import sys

import numpy as np
import pytest

import pyref
from openvm_b200 import _lib

P = pyref.P
R = pyref.BN_R


def limbs(states):
    return np.array([[w for x in st for w in pyref.bn_limbs(x)] for st in states], dtype=np.uint32)


def unlimbs(arr):
    return [[pyref.bn_from_limbs(row[8 * k:8 * k + 8]) for k in range(3)] for row in arr]


def random_states(n, seed):
    rng = np.random.default_rng(seed)
    out = [[int.from_bytes(rng.bytes(32), "little") % R for _ in range(3)] for _ in range(n)]
    out[0] = [0, 0, 0]
    out[1] = [R - 1, R - 1, R - 1]
    return out


def test_published_known_answer():
    # Poseidon2 reference implementation, bn256 instance, t = 3: permutation of (0, 1, 2)
    want = [0x0bb61d24daca55eebcb1929a82650f328134334da98ea4f847f760054f4a3033,
            0x303b6f7c86d043bfcbcc80214f26a30277a15d3f74ca654992defe7ff8d03570,
            0x1ed25194542b12eef8617361c3ba7c52e660b145994427cc86296242cf766ec8]
    assert pyref.bn_permute([0, 1, 2]) == want
    # leading round constants of zkhash poseidon2_instance_bn256.rs RC3 (recalled; the Grain LFSR regenerates the table)
    assert pyref.BN_EXT_INIT[0][0] == 0x1d066a255517b7fd8bddd3a93f7804ef7f8fcde48bb4c37a59a09a1a97052816
    assert pyref.BN_EXT_INIT[0][1] == 0x29daefb55f6f2dc6ac3f089cebcc6120b7c6fef31367b68eb7238547d32c1610


def test_oracle_permutation_matches_python(orc):
    states = random_states(24, 3) + [[0, 1, 2]]
    got = unlimbs(orc.bn254_permute(limbs(states)))
    assert got == [pyref.bn_permute(s) for s in states]


def test_product_host_permutation_matches_oracle(orc):
    lib = ctypes.CDLL(_lib.LIB_PATH)
    lib.b2s_host_poseidon2_bn254_permute.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
    states = limbs(random_states(200, 5) + [[0, 1, 2]])
    mine = states.copy()
    assert lib.b2s_host_poseidon2_bn254_permute(mine.ctypes.data, mine.shape[0]) == 0
    assert np.array_equal(mine, orc.bn254_permute(states))


def test_generated_header_is_current():
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tools"))
    import gen_bn254_constants as g
    assert open(g.path()).read() == g.render()


@pytest.fixture
def root_orc(orc):
    orc.set_hash(1)
    yield orc
    orc.set_hash(0)


def test_oracle_sponge_and_compression_match_python(root_orc):
    rng = np.random.default_rng(11)
    for n in (1, 7, 8, 9, 15, 16, 17, 31, 32, 33, 100):
        xs = rng.integers(0, P, size=n, dtype=np.uint32)
        xs[0] = P - 1
        assert pyref.bn_from_limbs(root_orc.hash_slice(xs)) == pyref.bn_hash_slice([int(v) for v in xs]), n
    # packed words of eight maximal values exceed the modulus: the sum is reduced, not truncated
    xs = np.full(16, P - 1, dtype=np.uint32)
    assert pyref.bn_from_limbs(root_orc.hash_slice(xs)) == pyref.bn_hash_slice([P - 1] * 16)
    a, b = random_states(2, 13)[1], random_states(3, 14)[2]
    for l, r in ((a[0], b[0]), (0, 0), (R - 1, 1)):
        got = root_orc.compress(np.array(pyref.bn_limbs(l), dtype=np.uint32), np.array(pyref.bn_limbs(r), dtype=np.uint32))
        assert pyref.bn_from_limbs(got) == pyref.bn_compress(l, r)


def test_oracle_challenger_matches_python(root_orc):
    rng = np.random.default_rng(17)
    digests = limbs([[int.from_bytes(rng.bytes(32), "little") % R, 0, 0] for _ in range(4)])[:, :8]
    ops, args = [], []
    ch = pyref.BnChallenger()
    want = []
    script = [0] * 3 + [2] + [0] * 3 + [2, 2] + [1] + [0] * 7 + [2] * 11 + [1, 1, 2] + [0] * 6 + [2, 0, 2]
    for k, op in enumerate(script):
        ops.append(op)
        if op == 0:
            v = int(rng.integers(0, P))
            args.append(v)
            ch.observe(v)
        elif op == 1:
            i = k % 4
            args.append(i)
            ch.observe_digest(pyref.bn_from_limbs(digests[i]))
        else:
            args.append(0)
            want.append(ch.sample())
    got = root_orc.challenger_script(ops, args, digests)
    assert [int(v) for v in got] == want


def test_oracle_root_config_proof_verifies(orc):
    """The oracle's whole prover and verifier follow the switch: a proof made under the root configuration verifies under
    it, is rejected under the other, and differs from the inner configuration's proof of the same statement."""
    from openvm_b200 import airs
    from test_oracle_stark import interaction_traces
    fib, snd, rcv = airs.fibonacci_air(), airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)
    key = airs.key_blob([fib, snd, rcv], log_up_pow_bits=3)
    send, recv = interaction_traces(16, 32, 2, seed=4)
    t0, pv0 = airs.fibonacci_trace(16)
    none = np.zeros(0, dtype=np.uint32)
    inputs = [(0, [t0], pv0), (1, [send], none), (2, [recv], none)]
    inner = orc.stark_prove(key, inputs, 1, 3, 2)
    orc.set_hash(1)
    try:
        proof = orc.stark_prove(key, inputs, 1, 3, 2)
        assert orc.stark_verify(key, proof, 1, 3, 2) is None
        assert proof != inner and len(proof) == len(inner)
        assert orc.stark_verify(key, inner, 1, 3, 2) is not None
        bad = bytearray(proof)
        bad[len(bad) // 2] ^= 1
        assert orc.stark_verify(key, bytes(bad), 1, 3, 2) is not None
    finally:
        orc.set_hash(0)
    assert orc.stark_verify(key, proof, 1, 3, 2) is not None
    assert orc.stark_verify(key, inner, 1, 3, 2) is None
