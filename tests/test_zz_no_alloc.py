"""b2s_merkle_root_static (openvm_b200/csrc/noalloc.cu): the allocation-free Merkle root over the static __device__ scratch
array equals b2s_mmcs_commit's root on the same matrix, bit for bit, on two calls in a row (the second on a shorter
matrix, so stale digests of the first call sit in the scratch it reuses). Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
P = 2013265921
MAX_ROWS = 65536


def _static_root(lib, engine, dmat, out):
    st = lib.b2s_merkle_root_static(dmat.device_ptr, dmat.height, dmat.width, out.device_ptr, engine.stream)
    assert st == 0, (st, lib.b2s_last_error(None))
    # the kernel wrote canonical words; a raw (column-major, no conversion) download on the same stream reads them back
    return out.to_host(layout="col", repr="monty").reshape(-1)


def test_static_root_equals_mmcs_commit_twice(engine):
    from openvm_b200 import _lib
    lib = _lib.load()
    dev = engine.device()
    rng = np.random.default_rng(77)
    out = dev.alloc_matrix(8, 1)
    for h, w in [(1 << 12, 19), (1 << 7, 8), (MAX_ROWS, 3), (1, 5)]:
        m = rng.integers(0, P, size=(h, w), dtype=np.uint64).astype(np.uint32)
        dm = dev.transport_matrix_to_device(m)
        want, pcs = dev.mmcs_commit([dm])
        got = _static_root(lib, engine, dm, out)
        assert np.array_equal(got, want), (h, w, got, want)
        # and once more on the same input: the result does not depend on what the scratch held
        assert np.array_equal(_static_root(lib, engine, dm, out), want)
        pcs.free()
        dm.free()


def test_static_root_rejects_oversize(engine):
    from openvm_b200 import _lib
    lib = _lib.load()
    dev = engine.device()
    out = dev.alloc_matrix(8, 1)
    dm = dev.alloc_matrix(2 * MAX_ROWS, 1)
    st = lib.b2s_merkle_root_static(dm.device_ptr, 2 * MAX_ROWS, 1, out.device_ptr, engine.stream)
    assert st == -2
    assert b"B2S_STATIC_MAX_ROWS" in lib.b2s_last_error(None)
    assert lib.b2s_merkle_root_static(dm.device_ptr, 3 * 1024, 1, out.device_ptr, engine.stream) == -2   # not a power of two
    assert lib.b2s_merkle_root_static(dm.device_ptr, 1024, 4097, out.device_ptr, engine.stream) == -2
    assert lib.b2s_merkle_root_static(None, 1024, 1, out.device_ptr, engine.stream) == -1
