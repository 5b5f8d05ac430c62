"""GPU parity, commitment: Merkle roots, digest layers and openings equal the oracle's (the reference's own
cross-backend assertion is root equality: crates/vm/src/system/cuda/program.rs:159-173)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2013265921
rng = np.random.default_rng(11)


def rand(*shape):
    return rng.integers(0, P, size=shape, dtype=np.uint64).astype(np.uint32)


MMCS_CASES = [
    [(1, 1)],
    [(1, 20)],
    [(2, 8)],
    [(8, 3)],
    [(64, 8)],
    [(64, 9)],
    [(64, 16), (64, 1)],           # sponge chunk straddles two matrices
    [(16, 5), (4, 2), (4, 7), (1, 3)],
    [(2, 1), (256, 17), (128, 8), (8, 30)],
    [(4096, 33), (512, 70)],
]


@pytest.mark.parametrize("shapes", MMCS_CASES)
def test_mmcs_root_layers_openings(engine, orc, shapes):
    dev = engine.device()
    mats = [rand(h, w) for h, w in shapes]
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


@pytest.mark.parametrize("shapes,b", [
    ([(32, 2)], 1),                               # Fibonacci AIR shape (recursion/src/tests.rs:109)
    ([(1 << 10, 10)], 1),                         # two_adic_pcs.rs:631-751 single-batch shape
    ([(1 << 12, 33)], 2),
    ([(1 << 15, 2), (1 << 9, 36), (4, 5), (1, 7)], 1),
    ([(1 << 16, 21), (1 << 16, 3), (1 << 13, 40)], 1),
])
def test_commit_root(engine, engine_b2, orc, shapes, b):
    e = engine if b == 1 else engine_b2
    dev = e.device()
    mats = [rand(h, w) for h, w in shapes]
    dmats = [dev.transport_matrix_to_device(m) for m in mats]
    root, pcs = dev.commit(dmats)
    want_root, want_ldes = orc.commit(mats, b, want_lde=True)
    assert np.array_equal(root, want_root)
    for i in range(len(mats)):
        assert np.array_equal(pcs.matrix(i).to_host(), want_ldes[i])


def test_commit_wide_trace(engine, orc):
    # the benchmark's column count at an oracle-sized height
    m = orc.fill_synthetic(1 << 12, 256, 0xB200000000000001)
    dev = engine.device()
    d = dev.alloc_matrix(1 << 12, 256)
    dev.fill_synthetic(d, 0xB200000000000001)
    root, _ = dev.commit([d])
    assert np.array_equal(root, orc.commit([m], 1))


def test_commit_full_scale_properties(engine):
    # 2^22 x 64 (too slow for the oracle): commitment is deterministic, sensitive to a single cell, and an
    # opened row + path recomputes the root with device-side hashing of the same row (open/verify round trip
    # checked against the oracle's verifier on the opened data only)
    import oracle_binding
    orc = oracle_binding.load()
    dev = engine.device()
    h, w = 1 << 22, 64
    d = dev.alloc_matrix(h, w)
    dev.fill_synthetic(d, 5)
    r1, pcs = dev.commit([d])
    r2, _ = dev.commit([d])
    assert np.array_equal(r1, r2)
    idx = 2718281
    rows, proof = pcs.open_batch(idx)
    node = orc.hash_slice(rows[0])
    for lvl in range(pcs.log_max_height):
        sib = proof[lvl]
        node = orc.compress(sib, node) if (idx >> lvl) & 1 else orc.compress(node, sib)
    assert np.array_equal(node, r1)


@pytest.mark.parametrize("h,w,layout,repr_", [
    (1 << 10, 5, "row", "canonical"), (1 << 12, 40, "col", "canonical"), (1 << 16, 19, "row", "canonical"),
    (1 << 23, 24, "col", "canonical"),   # 3 column groups of 8: state carried across groups
    (1 << 22, 50, "row", "canonical"),   # 4 groups of 16, ragged last group, strided 2-D copies
    (1, 3, "row", "canonical"),
])
def test_commit_host_streamed_equals_upload_then_commit(engine, h, w, layout, repr_):
    """b2s_commit_host (pipelined upload + LDE + streaming sponge) == b2s_matrix_upload + b2s_commit: same root, same
    opening at a few indices (rows of the retained LDE and the authentication path)."""
    dev = engine.device()
    rng = np.random.default_rng(h ^ w)
    m = rng.integers(0, P, size=(h, w), dtype=np.uint32)
    src = m if layout == "row" else np.ascontiguousarray(m.T)
    r1, p1 = dev.commit_host(src, layout=layout, repr=repr_)
    dm = dev.transport_matrix_to_device(m)
    r2, p2 = dev.commit([dm])
    assert list(r1) == list(r2)
    for idx in (0, 1, (h << 1) - 1, (h << 1) // 3):
        (rows1, proof1), (rows2, proof2) = p1.open_batch(idx), p2.open_batch(idx)
        assert np.array_equal(rows1[0], rows2[0]) and np.array_equal(proof1, proof2)
    p1.free(); p2.free(); dm.free()


_LEAF_PATH_SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, {root!r})
from openvm_b200 import B2Engine, FriParameters
rng = np.random.default_rng(77)
e = B2Engine.new(FriParameters(log_blowup=1), 0)
dev = e.device()
out = []
for group in {shapes!r}:
    ms = [rng.integers(0, 2013265921, size=(h, w), dtype=np.uint64).astype(np.uint32) for h, w in group]
    root, pcs = dev.commit([dev.transport_matrix_to_device(m) for m in ms])
    out.append(" ".join(str(int(x)) for x in root))
e.close()
print("|".join(out))
"""


@pytest.mark.parametrize("knob", ["0", "4", "8", "off"])
def test_leaf_hash_code_paths_agree(orc, knob):
    """The tall leaf level is hashed by a persistent kernel whose warps run either the integer or the FP64 Poseidon2
    (merkle.cu, poseidon2_f64.cuh), whichever is free. Every path -- all FP64 (B2S_HASH_F64=0), the shipped 4 + 4 mix,
    all integer (8), the one-row-per-thread integer kernel (off) -- must give the oracle's root, on widths that end in a
    partial chunk and on trees with compress-and-inject levels."""
    import os, subprocess, sys
    # the last group joins shorter matrices below the leaves (compress-and-inject levels) and has plain compress levels
    shapes = [[(1 << 13, 19)], [(1 << 14, 8)], [(1 << 13, 1)], [(1 << 15, 70)], [(1 << 16, 9), (1 << 14, 13), (1 << 13, 3)]]
    root_dir = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, B2S_HASH_F64=knob, B2S_HASH_MIN_LOG="12")  # the product switches at 2^19 rows
    r = subprocess.run([sys.executable, "-c", _LEAF_PATH_SCRIPT.format(root=root_dir, shapes=shapes)], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = [[int(x) for x in part.split()] for part in r.stdout.strip().splitlines()[-1].split("|")]
    rng2 = np.random.default_rng(77)
    for group, g in zip(shapes, got):
        ms = [rng2.integers(0, P, size=(h, w), dtype=np.uint64).astype(np.uint32) for h, w in group]
        want_root = orc.commit(ms, 1)
        assert g == [int(x) for x in want_root], (knob, group)


# ---- the two-pipe persistent leaf kernel at its SHIPPED threshold (levels of >= 2^19 nodes, merkle.cu MIX_MIN_ROWS) ------
# No environment override: these are the code paths the full-size commit runs. Oracle cost: seconds.
@pytest.mark.parametrize("shapes", [
    [(1 << 19, 8)],                       # one full sponge block per row
    [(1 << 19, 19)],                      # ragged last block
    [(1 << 19, 40)],
    [(1 << 20, 3), (1 << 19, 9)],         # compress-and-inject level of 2^19 nodes (leaf_mixed_kernel<ColTable, PREV>)
    [(1 << 20, 11), (1 << 20, 6)],        # a sponge block straddling two matrices on a 2^20-row level
])
def test_mixed_pipe_leaf_kernel_at_shipped_threshold(engine, orc, shapes):
    import os
    assert "B2S_HASH_MIN_LOG" not in os.environ and "B2S_HASH_F64" not in os.environ
    dev = engine.device()
    big = np.random.default_rng(19)
    mats = [big.integers(0, P, size=(h, w), dtype=np.uint64).astype(np.uint32) for h, w in shapes]
    dmats = [dev.transport_matrix_to_device(m) for m in mats]
    root, pcs = dev.mmcs_commit(dmats)
    want_root, want_layers = orc.mmcs_commit(mats, want_layers=True)
    assert np.array_equal(root, want_root)
    hmax = max(h for h, _ in shapes)
    for level in (0, 1):                  # the leaf level and the level above it (the injected one in case 4)
        off = 0 if level == 0 else hmax
        n = hmax >> level
        assert np.array_equal(pcs.layer(level), want_layers[off:off + n]), f"layer {level}"
    rows, proof = pcs.open_batch(hmax - 3)
    ok, want_proof = orc.mmcs_open_verify(mats, hmax - 3)
    assert ok and np.array_equal(proof, want_proof)
    pcs.free()
    for d in dmats:
        d.free()


def test_streamed_host_commit_at_shipped_threshold(engine, orc):
    """b2s_commit_host on a trace tall enough for the persistent kernel's suspend / resume path (leaf_absorb over column
    groups: state planes between groups) -- root equals the oracle's commit."""
    dev = engine.device()
    h, w = 1 << 18, 24                    # LDE 2^19 rows; column groups of 8 force three absorb calls
    m = np.random.default_rng(23).integers(0, P, size=(h, w), dtype=np.uint64).astype(np.uint32)
    import os
    os.environ["B2S_COMMIT_HOST_GROUP_COLS"] = "8"
    try:
        root, pcs = dev.commit_host(m)
    finally:
        del os.environ["B2S_COMMIT_HOST_GROUP_COLS"]
    assert np.array_equal(root, orc.commit([m], 1))
    pcs.free()


def test_async_ntt_variant_gives_the_oracle_roots(orc):
    """B2S_NTT_ASYNC=1 routes the strided NTT passes through the cp.async double-buffered persistent kernel (an experiment
    kept in the tree with its measurement, profiles/r2zb_ntt_async.md): same roots as the oracle on heights whose plans have
    strided passes of 5 to 8 bits (2^14 ... 2^21 rows)."""
    import os, subprocess, sys
    shapes = [[(1 << 14, 5)], [(1 << 16, 9), (1 << 15, 3)], [(1 << 18, 4)], [(1 << 21, 2)]]
    root_dir = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, B2S_NTT_ASYNC="1")
    r = subprocess.run([sys.executable, "-c", _LEAF_PATH_SCRIPT.format(root=root_dir, shapes=shapes)], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = [[int(x) for x in part.split()] for part in r.stdout.strip().splitlines()[-1].split("|")]
    rng2 = np.random.default_rng(77)
    for group, g in zip(shapes, got):
        ms = [rng2.integers(0, P, size=(h, w), dtype=np.uint64).astype(np.uint32) for h, w in group]
        assert g == [int(x) for x in orc.commit(ms, 1)], group
