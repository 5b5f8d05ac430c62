"""The host side of the Fiat-Shamir transcript (openvm_b200/csrc/host_challenger.hpp) hashes with an AVX2 Poseidon2
(csrc/host_poseidon2_avx2.cpp). No device needed: the vector permutation is compared with the oracle's on the same states.
Reference: the challenger is DuplexChallenger<BabyBear, Poseidon2, 16, 8> (crates/stark-sdk/src/config/baby_bear_poseidon2.rs:46-60)."""
import ctypes

import numpy as np
import pytest

import pyref
from openvm_b200 import _lib

P = pyref.P
R = (1 << 32) % P
RINV = pow(R, P - 2, P)


def _avx2(states):
    lib = ctypes.CDLL(_lib.LIB_PATH)
    if not lib.b2_host_has_avx2():
        pytest.skip("no AVX2 on this host")
    lib.b2_poseidon2_permute_avx2.argtypes = [ctypes.c_void_p]
    lib.b2_poseidon2_permute_avx2.restype = None
    monty = (states.astype(np.uint64) * R % P).astype(np.uint32)
    for row in monty:
        lib.b2_poseidon2_permute_avx2(row.ctypes.data)
    return (monty.astype(np.uint64) * RINV % P).astype(np.uint32)


def test_avx2_permutation_matches_oracle(orc):
    rng = np.random.default_rng(7)
    states = rng.integers(0, P, size=(512, 16), dtype=np.uint32)
    # corners: zeros, all p - 1, one hot lanes, alternating extremes
    states[0] = 0
    states[1] = P - 1
    for i in range(16):
        states[2 + i] = 0
        states[2 + i, i] = P - 1
    states[18, ::2] = P - 1
    states[18, 1::2] = 1
    assert np.array_equal(_avx2(states), orc.permute(states))


def test_avx2_permutation_known_answer():
    # the width-16 test vector pyref pins the permutation with (tests/test_oracle.py uses the same one)
    x = np.arange(16, dtype=np.uint32).reshape(1, 16)
    assert np.array_equal(_avx2(x)[0], np.array(pyref.permute(list(range(16))), dtype=np.uint32))
