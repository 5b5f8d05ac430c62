"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so). TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by openvm_b200."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
P = 2013265921


def build():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])
    return LIB


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        vp = C.c_void_p
        lib.orc_two_adic_generator.restype = C.c_uint32
        lib.orc_two_adic_generator.argtypes = [C.c_uint32]
        lib.orc_eval_naive.restype = C.c_uint32
        lib.orc_eval_naive.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32]
        lib.orc_coset_lde.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, vp, C.c_int]
        lib.orc_dft.argtypes = [vp, C.c_uint32, C.c_int]
        lib.orc_poseidon2_permute.argtypes = [vp, C.c_size_t]
        lib.orc_hash_slice.argtypes = [vp, C.c_size_t, vp]
        lib.orc_compress.argtypes = [vp, vp, vp]
        lib.orc_mmcs_commit.argtypes = [vp, vp, vp, C.c_size_t, vp, vp, C.c_int]
        lib.orc_mmcs_open_verify.restype = C.c_int
        lib.orc_mmcs_open_verify.argtypes = [vp, vp, vp, C.c_size_t, C.c_uint64, vp, C.c_int]
        lib.orc_commit.argtypes = [vp, vp, vp, C.c_size_t, C.c_uint32, vp, vp, C.c_int]
        lib.orc_fill_synthetic.argtypes = [vp, C.c_uint64, C.c_uint64, C.c_uint64]
        lib.orc_ef_mul.argtypes = [vp, vp, vp]
        lib.orc_ef_inv.argtypes = [vp, vp]
        lib.orc_poseidon2_constants.argtypes = [vp]
        lib.orc_last_error.restype = C.c_char_p
        lib.orc_free.argtypes = [vp]
        lib.orc_stark_prove.restype = C.c_int
        lib.orc_stark_prove.argtypes = [vp, C.c_size_t, C.c_size_t, vp, vp, vp, vp, C.c_uint32, C.c_uint32, C.c_uint32,
                                        C.c_int, C.POINTER(vp), C.POINTER(C.c_size_t)]
        lib.orc_stark_verify.restype = C.c_int
        lib.orc_stark_verify.argtypes = [vp, C.c_size_t, vp, C.c_size_t, C.c_uint32, C.c_uint32, C.c_uint32]
        lib.orc_challenger_run.argtypes = [vp, C.c_size_t, vp, C.c_size_t]

    # ---- constants / field
    def poseidon2_constants(self):
        out = np.zeros(157, dtype=np.uint32)
        self.lib.orc_poseidon2_constants(out.ctypes.data)
        return {"ext_init": out[:64].reshape(4, 16), "partial": out[64:77], "ext_term": out[77:141].reshape(4, 16),
                "diag": out[141:157]}

    def two_adic_generator(self, bits):
        return int(self.lib.orc_two_adic_generator(bits))

    def ef_mul(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.uint32); b = np.ascontiguousarray(b, dtype=np.uint32)
        out = np.zeros(4, dtype=np.uint32)
        self.lib.orc_ef_mul(a.ctypes.data, b.ctypes.data, out.ctypes.data)
        return out

    def ef_inv(self, a):
        a = np.ascontiguousarray(a, dtype=np.uint32)
        out = np.zeros(4, dtype=np.uint32)
        self.lib.orc_ef_inv(a.ctypes.data, out.ctypes.data)
        return out

    # ---- poseidon2
    def permute(self, states):
        a = np.ascontiguousarray(states, dtype=np.uint32).reshape(-1, 16).copy()
        self.lib.orc_poseidon2_permute(a.ctypes.data, a.shape[0])
        return a

    def hash_slice(self, x):
        x = np.ascontiguousarray(x, dtype=np.uint32)
        out = np.zeros(8, dtype=np.uint32)
        self.lib.orc_hash_slice(x.ctypes.data, x.size, out.ctypes.data)
        return out

    def compress(self, l, r):
        l = np.ascontiguousarray(l, dtype=np.uint32); r = np.ascontiguousarray(r, dtype=np.uint32)
        out = np.zeros(8, dtype=np.uint32)
        self.lib.orc_compress(l.ctypes.data, r.ctypes.data, out.ctypes.data)
        return out

    # ---- dft / lde (matrices are (height, width) row-major numpy arrays of canonical values)
    def dft(self, col, inverse=False):
        a = np.ascontiguousarray(col, dtype=np.uint32).copy()
        logn = int(a.size).bit_length() - 1
        self.lib.orc_dft(a.ctypes.data, logn, 1 if inverse else 0)
        return a

    def eval_naive(self, evals, shift_in, x):
        a = np.ascontiguousarray(evals, dtype=np.uint32)
        return int(self.lib.orc_eval_naive(a.ctypes.data, int(a.size).bit_length() - 1, shift_in, x))

    def coset_lde(self, mat, log_blowup, shift_in=1, threads=8):
        h, w = mat.shape
        cm = np.ascontiguousarray(mat.T, dtype=np.uint32)
        logn = h.bit_length() - 1
        out = np.zeros((w, h << log_blowup), dtype=np.uint32)
        self.lib.orc_coset_lde(cm.ctypes.data, logn, w, log_blowup, shift_in, out.ctypes.data, threads)
        return np.ascontiguousarray(out.T)

    def _mat_args(self, mats):
        cms = [np.ascontiguousarray(m.T, dtype=np.uint32) for m in mats]
        ptrs = (C.c_void_p * len(mats))(*[c.ctypes.data for c in cms])
        heights = np.array([m.shape[0] for m in mats], dtype=np.uint64)
        widths = np.array([m.shape[1] for m in mats], dtype=np.uint64)
        return cms, ptrs, heights, widths

    def mmcs_commit(self, mats, threads=8, want_layers=False):
        cms, ptrs, heights, widths = self._mat_args(mats)
        root = np.zeros(8, dtype=np.uint32)
        layers = None
        if want_layers:
            hmax = int(heights.max())
            layers = np.zeros((2 * hmax - 1, 8), dtype=np.uint32)
        self.lib.orc_mmcs_commit(ptrs, heights.ctypes.data, widths.ctypes.data, len(mats), root.ctypes.data,
                                 layers.ctypes.data if want_layers else None, threads)
        return (root, layers) if want_layers else root

    def mmcs_open_verify(self, mats, index, corrupt=False):
        cms, ptrs, heights, widths = self._mat_args(mats)
        logh = int(heights.max()).bit_length() - 1
        proof = np.zeros((logh, 8), dtype=np.uint32)
        ok = self.lib.orc_mmcs_open_verify(ptrs, heights.ctypes.data, widths.ctypes.data, len(mats), index,
                                           proof.ctypes.data, 1 if corrupt else 0)
        return bool(ok), proof

    def commit(self, mats, log_blowup, threads=8, want_lde=False):
        cms, ptrs, heights, widths = self._mat_args(mats)
        logs = np.array([int(h).bit_length() - 1 for h in heights], dtype=np.uint64)
        root = np.zeros(8, dtype=np.uint32)
        ldes, lde_ptrs = None, None
        if want_lde:
            ldes = [np.zeros((m.shape[1], m.shape[0] << log_blowup), dtype=np.uint32) for m in mats]
            lde_ptrs = (C.c_void_p * len(mats))(*[a.ctypes.data for a in ldes])
        self.lib.orc_commit(ptrs, logs.ctypes.data, widths.ctypes.data, len(mats), log_blowup, root.ctypes.data,
                            lde_ptrs, threads)
        if want_lde:
            return root, [np.ascontiguousarray(a.T) for a in ldes]
        return root

    # ---- STARK
    def stark_prove(self, key_words, airs, log_blowup, num_queries, pow_bits, threads=8):
        """airs: list of (air_id, [matrices (height, width) uint32: cached parts then common], public_values)."""
        key = np.ascontiguousarray(key_words, dtype=np.uint32)
        ids = np.array([a[0] for a in airs], dtype=np.uint32)
        heights = np.array([a[1][0].shape[0] for a in airs], dtype=np.uint64)
        cols = [np.ascontiguousarray(m.T, dtype=np.uint32) for a in airs for m in a[1]]
        mat_ptrs = (C.c_void_p * max(1, len(cols)))(*[c.ctypes.data for c in cols])
        pvs = [np.ascontiguousarray(a[2], dtype=np.uint32) for a in airs]
        pv_ptrs = (C.c_void_p * len(airs))(*[p.ctypes.data for p in pvs])
        out, n = C.c_void_p(), C.c_size_t()
        rc = self.lib.orc_stark_prove(key.ctypes.data, key.size, len(airs), ids.ctypes.data, mat_ptrs, heights.ctypes.data,
                                      pv_ptrs, log_blowup, num_queries, pow_bits, threads, C.byref(out), C.byref(n))
        if rc != 0:
            raise RuntimeError("oracle prove: " + self.lib.orc_last_error().decode())
        data = C.string_at(out.value, n.value)
        self.lib.orc_free(out)
        return data

    def stark_verify(self, key_words, proof, log_blowup, num_queries, pow_bits):
        """Returns None when accepted, else the name of the verification error."""
        key = np.ascontiguousarray(key_words, dtype=np.uint32)
        buf = (C.c_uint8 * len(proof)).from_buffer_copy(proof)
        rc = self.lib.orc_stark_verify(key.ctypes.data, key.size, buf, len(proof), log_blowup, num_queries, pow_bits)
        return None if rc == 0 else self.lib.orc_last_error().decode()

    def proof_hint_stream(self, proof):
        """Proof bytes -> list of vectors (lists of canonical words): oracle/hints.hpp."""
        buf = (C.c_uint8 * len(proof)).from_buffer_copy(proof)
        words, n = C.POINTER(C.c_uint32)(), C.c_size_t()
        self.lib.orc_proof_hint_stream.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.POINTER(C.c_uint32)), C.POINTER(C.c_size_t)]
        rc = self.lib.orc_proof_hint_stream(buf, len(proof), C.byref(words), C.byref(n))
        if rc != 0:
            raise ValueError(self.lib.orc_last_error().decode())
        flat = np.ctypeslib.as_array(words, shape=(n.value,)).copy()
        self.lib.orc_free(words)
        return unflatten_stream(flat)

    # ---- root configuration (oracle/bn254.hpp)
    def set_hash(self, kind):
        self.lib.orc_set_hash(int(kind))

    def bn254_permute(self, states):
        a = np.ascontiguousarray(states, dtype=np.uint32).reshape(-1, 24).copy()
        self.lib.orc_bn254_permute(C.c_void_p(a.ctypes.data), C.c_size_t(a.shape[0]))
        return a

    def challenger_script(self, ops, args, digests):
        ops = np.ascontiguousarray(ops, dtype=np.uint32); args = np.ascontiguousarray(args, dtype=np.uint32)
        digests = np.ascontiguousarray(digests, dtype=np.uint32).reshape(-1)
        out = np.zeros(max(1, int((ops == 2).sum())), dtype=np.uint32)
        self.lib.orc_challenger_script(C.c_void_p(ops.ctypes.data), C.c_void_p(args.ctypes.data), C.c_size_t(ops.size),
                                       C.c_void_p(digests.ctypes.data), C.c_void_p(out.ctypes.data))
        return out[:int((ops == 2).sum())]

    def proof_witness_stream(self, proof):
        """Root proof bytes -> (vars [n, 8], felts [n], exts [n, 4]): oracle/witness.hpp."""
        buf = (C.c_uint8 * len(proof)).from_buffer_copy(proof)
        words, n = C.POINTER(C.c_uint32)(), C.c_size_t()
        self.lib.orc_proof_witness_stream.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.POINTER(C.c_uint32)), C.POINTER(C.c_size_t)]
        rc = self.lib.orc_proof_witness_stream(buf, len(proof), C.byref(words), C.byref(n))
        if rc != 0:
            raise RuntimeError(self.lib.orc_last_error().decode())
        flat = np.ctypeslib.as_array(words, shape=(n.value,)).copy()
        self.lib.orc_free(words)
        nv, nf, ne = (int(x) for x in flat[:3])
        a, b = 3 + 8 * nv, 3 + 8 * nv + nf
        return flat[3:a].reshape(nv, 8), flat[a:b], flat[b:b + 4 * ne].reshape(ne, 4)

    def challenger_run(self, observed, n_samples):
        obs = np.ascontiguousarray(observed, dtype=np.uint32)
        out = np.zeros(n_samples, dtype=np.uint32)
        self.lib.orc_challenger_run(obs.ctypes.data, obs.size, out.ctypes.data, n_samples)
        return out

    def fast_commit(self, log_h, width, log_blowup, matrix=None, seed=0, threads=8):
        """orc_fast_commit (oracle/fast_commit.cpp): AVX2 / Montgomery / multi-threaded commit of one matrix -- either
        `matrix` ((height, width) canonical) or the synthetic trace of `seed` generated column by column.
        Returns (root, {"gen_s", "lde_s", "hash_s"})."""
        self.lib.orc_fast_commit.restype = C.c_int
        self.lib.orc_fast_commit.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32, C.c_int, C.c_void_p,
                                             C.c_void_p]
        cm = None
        if matrix is not None:
            cm = np.ascontiguousarray(np.asarray(matrix, dtype=np.uint32).T)
            assert cm.shape == (width, 1 << log_h)
        root = np.zeros(8, dtype=np.uint32)
        times = np.zeros(3, dtype=np.float64)
        rc = self.lib.orc_fast_commit(cm.ctypes.data if cm is not None else None, seed, log_h, width, log_blowup, threads,
                                      root.ctypes.data, times.ctypes.data)
        if rc != 0:
            raise ValueError("orc_fast_commit: bad arguments")
        return root, {"gen_s": float(times[0]), "lde_s": float(times[1]), "hash_s": float(times[2])}

    def fill_synthetic(self, height, width, seed):
        out = np.zeros((width, height), dtype=np.uint32)
        self.lib.orc_fill_synthetic(out.ctypes.data, height, width, seed)
        return np.ascontiguousarray(out.T)


_oracle = None


def load():
    global _oracle
    if _oracle is None:
        build()
        _oracle = Oracle(C.CDLL(LIB))
    return _oracle


def unflatten_stream(flat):
    """n_vectors, then per vector its length and words -> list of lists"""
    flat = [int(x) for x in flat]
    out, i = [], 1
    for _ in range(flat[0]):
        ln = flat[i]
        out.append(flat[i + 1:i + 1 + ln])
        i += 1 + ln
    assert i == len(flat)
    return out
