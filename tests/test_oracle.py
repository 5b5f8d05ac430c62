"""The C++ oracle against an independent pure-Python restatement and against definitional checks."""
import numpy as np
import pytest

import pyref

P = pyref.P
rng = np.random.default_rng(20261104)


def rand(*shape):
    return rng.integers(0, P, size=shape, dtype=np.uint64).astype(np.uint32)


def test_permutation_matches_python(orc):
    st = rand(5, 16)
    st[0] = 0
    st[1] = np.arange(16)
    got = orc.permute(st)
    for i in range(5):
        assert [int(x) for x in got[i]] == pyref.permute([int(x) for x in st[i]])


def test_permutation_is_not_identity_and_is_deterministic(orc):
    z = orc.permute(np.zeros((1, 16), dtype=np.uint32))
    assert z.any()
    assert np.array_equal(z, orc.permute(np.zeros((1, 16), dtype=np.uint32)))


def test_sponge_overwrite_semantics(orc):
    # PaddingFreeSponge: short last chunk overwrites only its own lanes (README.md:9-17)
    for n in (0, 1, 7, 8, 9, 16, 23):
        x = rand(n)
        assert [int(v) for v in orc.hash_slice(x)] == pyref.hash_slice([int(v) for v in x])
    # no padding: [a] and [a, 0] differ only because lane 1 is overwritten with 0 == initial 0 -> equal digests
    a = rand(1)
    assert np.array_equal(orc.hash_slice(a), orc.hash_slice(np.concatenate([a, np.zeros(1, dtype=np.uint32)])))


def test_compress(orc):
    l, r = rand(8), rand(8)
    assert [int(v) for v in orc.compress(l, r)] == pyref.compress([int(v) for v in l], [int(v) for v in r])


def test_ext_field(orc):
    a, b = rand(4), rand(4)
    ab = orc.ef_mul(a, b)
    # schoolbook with x^4 = 11
    t = [0] * 7
    for i in range(4):
        for j in range(4):
            t[i + j] += int(a[i]) * int(b[j])
    want = [(t[0] + 11 * t[4]) % P, (t[1] + 11 * t[5]) % P, (t[2] + 11 * t[6]) % P, t[3] % P]
    assert [int(x) for x in ab] == want
    one = orc.ef_mul(a, orc.ef_inv(a))
    assert [int(x) for x in one] == [1, 0, 0, 0]


def test_dft_roundtrip_and_definition(orc):
    for logn in (0, 1, 3, 6):
        n = 1 << logn
        x = rand(n)
        X = orc.dft(x)
        g = pyref.two_adic_generator(logn)
        for k in (0, n // 2, n - 1):
            assert int(X[k]) == sum(int(x[i]) * pow(g, i * k, P) for i in range(n)) % P
        assert np.array_equal(orc.dft(X, inverse=True), x)


def test_coset_lde_matches_naive(orc):
    for logn, b, shift in ((0, 1, 1), (1, 1, 1), (3, 1, 1), (4, 2, 1), (3, 2, 31 * 7 % P), (5, 3, 1)):
        m = rand(1 << logn, 3)
        got = orc.coset_lde(m, b, shift)
        for c in range(3):
            assert [int(v) for v in got[:, c]] == pyref.coset_lde_naive([int(v) for v in m[:, c]], b, shift)


def test_lde_leaf_index_convention(orc):
    # two_adic_pcs.rs:178-186: row i of the LDE is the evaluation at GENERATOR * w^{bitrev(i)}
    logn, b = 4, 1
    col = rand(1 << logn, 1)
    lde = orc.coset_lde(col, b)
    w = pyref.two_adic_generator(logn + b)
    for i in (0, 1, 5, 31):
        br = int(format(i, "05b")[::-1], 2)
        x = 31 * pow(w, br, P) % P
        assert int(lde[i, 0]) == orc.eval_naive(col[:, 0], 1, x)


def test_mmcs_matches_python_mixed_heights(orc):
    cases = [
        [(8, 3)],
        [(8, 9), (8, 1)],
        [(4, 2), (16, 5), (4, 7), (1, 3)],
        [(2, 1), (16, 17), (8, 8)],
        [(1, 5)],
    ]
    for shapes in cases:
        mats = [rand(h, w) for h, w in shapes]
        want = pyref.mmcs_root([[[int(v) for v in row] for row in m] for m in mats])
        assert [int(v) for v in orc.mmcs_commit(mats)] == want


def test_mmcs_open_verify(orc):
    mats = [rand(4, 2), rand(16, 5), rand(4, 7), rand(8, 3)]
    for idx in (0, 5, 15):
        ok, proof = orc.mmcs_open_verify(mats, idx)
        assert ok and proof.shape == (4, 8)
        bad, _ = orc.mmcs_open_verify(mats, idx, corrupt=True)
        assert not bad


def test_commit_is_lde_then_mmcs(orc):
    mats = [rand(8, 3), rand(32, 2)]
    root, ldes = orc.commit(mats, 1, want_lde=True)
    assert np.array_equal(ldes[0], orc.coset_lde(mats[0], 1))
    assert np.array_equal(root, orc.mmcs_commit(ldes))


def test_synthetic_fill_is_deterministic(orc):
    a = orc.fill_synthetic(16, 3, 0xB200000000000001)
    assert a.shape == (16, 3) and (a < P).all()
    assert np.array_equal(a, orc.fill_synthetic(16, 3, 0xB200000000000001))
    assert not np.array_equal(a, orc.fill_synthetic(16, 3, 2))


@pytest.mark.parametrize("log_h,w,b", [(0, 3, 1), (1, 8, 1), (2, 9, 1), (3, 1, 2), (5, 17, 1), (10, 19, 1), (12, 40, 2),
                                       (14, 64, 1), (16, 24, 1), (13, 100, 2)])
def test_fast_commit_equals_plain_oracle(orc, log_h, w, b):
    """oracle/fast_commit.cpp (AVX2, Montgomery, multi-threaded; bench.py's CPU baseline and the checker of the full-size root)
    against the plain oracle's commit (canonical % arithmetic, pinned by the golden vectors): same root for tiny, ragged and
    multi-batch shapes, for any thread count."""
    rng = np.random.default_rng(log_h * 131 + w)
    m = rng.integers(0, 2013265921, size=(1 << log_h, w), dtype=np.uint64).astype(np.uint32)
    want = orc.commit([m], b)
    for threads in (1, 3, 8):
        got, times = orc.fast_commit(log_h, w, b, matrix=m, threads=threads)
        assert np.array_equal(got, want), (threads, got, want)
        assert times["lde_s"] >= 0 and times["hash_s"] >= 0


def test_fast_commit_generates_the_synthetic_trace(orc):
    m = orc.fill_synthetic(1 << 11, 16, 0xB200000000000001)
    assert np.array_equal(orc.commit([m], 1), orc.fast_commit(11, 16, 1, seed=0xB200000000000001, threads=4)[0])
    with pytest.raises(ValueError):
        orc.fast_commit(27, 1, 1, seed=1)
