"""Host-side mirror of the reference's prover interface for the STARK hot path.

Names and argument meaning follow the (external) openvm-stark-backend traits as they are used in-tree:

    E::new(fri_params)                              crates/sdk/src/prover/vm/mod.rs:22
    engine.fri_params()                             crates/sdk/src/prover/app.rs:139
    engine.device()                                 crates/vm/src/arch/vm.rs:382
    device.transport_matrix_to_device(matrix)       crates/vm/src/arch/vm.rs:731-734
    device.commit(&[matrix]) -> (Com, PcsData)      crates/vm/src/arch/vm.rs:735
    device.transport_matrix_from_device_to_host     crates/vm/src/arch/vm.rs:1307-1312

Everything here forwards to libb2stark.so through ctypes; numpy arrays are only host buffers.
"""
import ctypes as C
import weakref
from dataclasses import dataclass

import numpy as np

from . import _lib

P = 2013265921  # BabyBear


class B2StarkError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"b2stark status {status}: {message}")
        self.status = status


class VerificationError(Exception):
    """openvm_stark_backend::verifier::VerificationError (variant name in the message)."""


def verify_host(key_words, proof, fri_params, hash_kind=0):
    """StarkEngine::verify without a device (b2s_verify_host): the verifier is host code. Returns None, or raises
    VerificationError carrying the reference's error variant name."""
    lib = _lib.load()
    key = np.ascontiguousarray(key_words, dtype=np.uint32)
    p = _lib.FriParams(fri_params.log_blowup, fri_params.log_final_poly_len, fri_params.num_queries, fri_params.proof_of_work_bits)
    proof = bytes(proof)
    st = lib.b2s_verify_host(C.byref(p), int(hash_kind), key.ctypes.data, key.size, proof, len(proof))
    if st == _lib.ERR_VERIFY_FAILED:
        raise VerificationError((lib.b2s_last_error(None) or b"").decode())
    if st != 0:
        raise B2StarkError(st, (lib.b2s_last_error(None) or b"").decode())


def library_path():
    return _lib.LIB_PATH


@dataclass(frozen=True)
class FriParameters:
    """openvm_stark_sdk::config::FriParameters (defaults: crates/sdk/src/config/mod.rs:15-18)."""
    log_blowup: int = 1
    log_final_poly_len: int = 0
    num_queries: int = 100
    proof_of_work_bits: int = 16

    @staticmethod
    def standard_with_100_bits_conjectured_security(log_blowup):
        # docs/vocs/docs/pages/specs/security/security-model.mdx:47-53
        q = {1: 100, 2: 44, 3: 30, 4: 23}[log_blowup]
        return FriParameters(log_blowup, 0, q, 16)

    @staticmethod
    def new_for_testing(log_blowup):
        return FriParameters(log_blowup, 0, 2, 0)


class DeviceMatrix:
    """Column-major BabyBear matrix in HBM (Montgomery form); mirrors DeviceMatrix<F>{buffer,height,width}."""

    def __init__(self, engine, handle, borrowed=False):
        self._engine = engine
        self._h = handle
        self._borrowed = borrowed
        engine._track(self)

    @property
    def height(self):
        return int(self._engine._lib.b2s_matrix_height(self._h))

    @property
    def width(self):
        return int(self._engine._lib.b2s_matrix_width(self._h))

    @property
    def device_ptr(self):
        return int(self._engine._lib.b2s_matrix_device_ptr(self._h) or 0)

    def to_host(self, layout="row", repr="canonical"):
        """Returns a (height, width) numpy uint32 array of canonical values (row-major by default)."""
        return self._engine.device().transport_matrix_from_device_to_host(self, layout, repr)

    def free(self):
        if self._h and not self._borrowed:
            self._engine._lib.b2s_matrix_free(self._h)
        self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PcsData:
    """Merkle tree + retained LDE matrices of one commitment (cpu::PcsData / GPU PcsData analogue;
    crates/vm/src/system/program/trace.rs:74-82)."""

    def __init__(self, engine, handle):
        self._engine = engine
        self._h = handle
        engine._track(self)

    @property
    def log_max_height(self):
        return int(self._engine._lib.b2s_pcs_log_max_height(self._h))

    def num_matrices(self):
        return int(self._engine._lib.b2s_pcs_num_matrices(self._h))

    def matrix(self, i):
        h = self._engine._lib.b2s_pcs_matrix(self._h, i)
        if not h:
            raise IndexError(i)
        m = DeviceMatrix(self._engine, h, borrowed=True)
        m._keepalive = self
        return m

    def open_batch(self, index):
        """Mmcs::open_batch: (rows in caller order, sibling digests bottom-up), canonical values."""
        e = self._engine
        widths = [self.matrix(i).width for i in range(self.num_matrices())]
        rows = np.zeros(sum(widths), dtype=np.uint32)
        proof = np.zeros((self.log_max_height, 8), dtype=np.uint32)
        e._check(e._lib.b2s_pcs_open(e._h, self._h, index, rows.ctypes.data, proof.ctypes.data))
        out, off = [], 0
        for w in widths:
            out.append(rows[off:off + w].copy())
            off += w
        return out, proof

    def layer(self, level):
        e = self._engine
        n = (1 << self.log_max_height) >> level
        out = np.zeros((n, 8), dtype=np.uint32)
        e._check(e._lib.b2s_pcs_layer(e._h, self._h, level, out.ctypes.data))
        return out

    def free(self):
        if self._h:
            self._engine._lib.b2s_pcs_free(self._h)
        self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class ProvingKey:
    """Device-resident proving key (transport_pk_to_device; crates/vm/src/arch/vm.rs:382)."""

    def __init__(self, engine, handle):
        self._engine = engine
        self._h = handle
        engine._track(self)

    def free(self):
        if self._h:
            self._engine._lib.b2s_pk_free(self._h)
        self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


@dataclass
class AirProvingContext:
    """AirProvingContext{cached_mains, common_main, public_values} for one AIR (crates/vm/src/arch/vm.rs:588-599).
    cached_mains are PcsData objects returned by device.commit([trace]); cached_traces the device matrices they
    were committed from (CommittedTraceData carries both), needed only by AIRs with interactions."""
    air_id: int
    common_main: object = None
    public_values: object = ()
    cached_mains: tuple = ()
    cached_traces: tuple = ()


class B2Device:
    """ProverDevice: TraceCommitter + DeviceDataTransporter (QuotientCommitter / OpeningProver: see prover)."""

    def __init__(self, engine):
        self._e = engine

    def transport_matrix_to_device(self, matrix, layout="row", repr="canonical"):
        """matrix: (height, width) uint32 array. layout 'row' = RowMajorMatrix memory, 'col' = column-major
        (then `matrix` is the (width, height) array of columns)."""
        e = self._e
        a = np.ascontiguousarray(matrix, dtype=np.uint32)
        if a.ndim != 2:
            raise ValueError("matrix must be 2-D")
        if layout == "row":
            height, width = a.shape
        else:
            width, height = a.shape
        h = C.c_void_p()
        e._check(e._lib.b2s_matrix_upload(e._h, a.ctypes.data, height, width, 0 if layout == "row" else 1,
                                          0 if repr == "canonical" else 1, C.byref(h)))
        return DeviceMatrix(e, h)

    def upload_raw(self, host_ptr, height, width, layout="col", repr="monty"):
        """Upload from a raw host pointer (e.g. pinned torch tensor data_ptr())."""
        e = self._e
        h = C.c_void_p()
        e._check(e._lib.b2s_matrix_upload(e._h, host_ptr, height, width, 0 if layout == "row" else 1,
                                          0 if repr == "canonical" else 1, C.byref(h)))
        return DeviceMatrix(e, h)

    def adopt_device_matrix(self, device_ptr, height, width):
        e = self._e
        h = C.c_void_p()
        e._check(e._lib.b2s_matrix_adopt_device(e._h, device_ptr, height, width, C.byref(h)))
        return DeviceMatrix(e, h)

    def alloc_matrix(self, height, width):
        e = self._e
        h = C.c_void_p()
        e._check(e._lib.b2s_matrix_alloc(e._h, height, width, C.byref(h)))
        return DeviceMatrix(e, h)

    def fill_synthetic(self, matrix, seed):
        e = self._e
        e._check(e._lib.b2s_matrix_fill_synthetic(e._h, matrix._h, seed))

    def fill_chain(self, matrix, seed):
        """Satisfying trace of airs.chain_air written on the device (b2s_matrix_fill_chain); returns the two public values."""
        e = self._e
        pv = np.zeros(2, dtype=np.uint32)
        e._check(e._lib.b2s_matrix_fill_chain(e._h, matrix._h, seed, pv.ctypes.data_as(_lib.u32p)))
        return pv

    def transport_matrix_from_device_to_host(self, matrix, layout="row", repr="canonical"):
        e = self._e
        h, w = matrix.height, matrix.width
        out = np.zeros((h, w) if layout == "row" else (w, h), dtype=np.uint32)
        e._check(e._lib.b2s_matrix_download(e._h, matrix._h, out.ctypes.data, 0 if layout == "row" else 1,
                                            0 if repr == "canonical" else 1))
        return out

    def download_raw(self, matrix, host_ptr, layout="col", repr="monty"):
        """Download into a raw host pointer with room for height*width words."""
        e = self._e
        e._check(e._lib.b2s_matrix_download(e._h, matrix._h, host_ptr, 0 if layout == "row" else 1,
                                            0 if repr == "canonical" else 1))

    def _handles(self, mats):
        arr = (C.c_void_p * len(mats))(*[m._h for m in mats])
        return arr

    def commit(self, matrices):
        """TraceCommitter::commit(&[matrix]) -> (commitment as 8 canonical words, PcsData)."""
        e = self._e
        root = np.zeros(8, dtype=np.uint32)
        out = C.c_void_p()
        e._check(e._lib.b2s_commit(e._h, self._handles(matrices), len(matrices),
                                   root.ctypes.data_as(_lib.u32p), C.byref(out)))
        return root, PcsData(e, out)

    def commit_host(self, matrix=None, layout="row", repr="canonical", host_ptr=None, height=None, width=None):
        """transport_matrix_to_device + commit of one tall host trace as a pipelined call (b2s_commit_host): the upload
        of column groups overlaps the LDE and hashing of the previous group. Either `matrix` (a (height, width) array,
        or (width, height) for layout 'col') or a raw `host_ptr` with height / width (e.g. pinned torch memory)."""
        e = self._e
        if matrix is not None:
            a = np.ascontiguousarray(matrix, dtype=np.uint32)
            height, width = a.shape if layout == "row" else a.shape[::-1]
            host_ptr = a.ctypes.data
        root = np.zeros(8, dtype=np.uint32)
        out = C.c_void_p()
        e._check(e._lib.b2s_commit_host(e._h, host_ptr, height, width, 0 if layout == "row" else 1,
                                        0 if repr == "canonical" else 1, root.ctypes.data, C.byref(out)))
        return root, PcsData(e, out)

    def coset_lde(self, matrix, log_blowup, shift=1):
        e = self._e
        out = C.c_void_p()
        e._check(e._lib.b2s_coset_lde(e._h, matrix._h, log_blowup, shift, C.byref(out)))
        return DeviceMatrix(e, out)

    def mmcs_commit(self, matrices):
        e = self._e
        root = np.zeros(8, dtype=np.uint32)
        out = C.c_void_p()
        e._check(e._lib.b2s_mmcs_commit(e._h, self._handles(matrices), len(matrices),
                                        root.ctypes.data_as(_lib.u32p), C.byref(out)))
        pcs = PcsData(e, out)
        pcs._keepalive = list(matrices)
        return root, pcs

    def transport_pk_to_device(self, key_words):
        """key_words: uint32 key blob (include/b2stark.h: b2s_pk_load)."""
        e = self._e
        a = np.ascontiguousarray(key_words, dtype=np.uint32)
        h = C.c_void_p()
        e._check(e._lib.b2s_pk_load(e._h, a.ctypes.data, a.size, C.byref(h)))
        return ProvingKey(e, h)

    def poseidon2_permute(self, states):
        e = self._e
        a = np.ascontiguousarray(states, dtype=np.uint32).reshape(-1, 16).copy()
        e._check(e._lib.b2s_poseidon2_permute(e._h, a.ctypes.data, a.shape[0]))
        return a

    def poseidon2_bn254_permute(self, states):
        """Poseidon2 over the BN254 scalar field, width 3: states of 3 x 8 canonical 32-bit limbs."""
        e = self._e
        a = np.ascontiguousarray(states, dtype=np.uint32).reshape(-1, 24).copy()
        e._check(e._lib.b2s_poseidon2_bn254_permute(e._h, a.ctypes.data, a.shape[0]))
        return a


HASH_POSEIDON2_BABYBEAR = 0   # BabyBearPoseidon2Config
HASH_POSEIDON2_BN254 = 1      # BabyBearPoseidon2RootConfig (crates/continuations/src/lib.rs:17)


class B2Engine:
    """StarkFriEngine<SC = BabyBearPoseidon2Config> on one B200 (set_hash selects the root configuration's hash)."""

    def __init__(self, fri_params=FriParameters(), cuda_device=0):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self._live = weakref.WeakSet()
        p = _lib.FriParams(fri_params.log_blowup, fri_params.log_final_poly_len, fri_params.num_queries,
                           fri_params.proof_of_work_bits)
        st = self._lib.b2s_engine_new(C.byref(p), cuda_device, C.byref(self._h))
        if st != 0:
            msg = self._lib.b2s_last_error(None)
            self._h = None
            raise B2StarkError(st, msg.decode() if msg else "")
        self._params = fri_params
        self._device = B2Device(self)

    @classmethod
    def new(cls, fri_params, cuda_device=0):
        return cls(fri_params, cuda_device)

    def fri_params(self):
        return self._params

    def set_hash(self, kind):
        self._check(self._lib.b2s_engine_set_hash(self._h, int(kind)))

    def hash(self):
        return int(self._lib.b2s_engine_hash(self._h))

    def prove(self, pk, ctx):
        """StarkEngine::prove(&pk, ctx) -> proof bytes in the reference wire format (crates/sdk/src/codec.rs).
        ctx: list of AirProvingContext in strictly increasing air_id."""
        n = len(ctx)
        arr = (_lib.AirCtx * n)()
        keep = []
        for i, a in enumerate(ctx):
            pv = np.ascontiguousarray(a.public_values, dtype=np.uint32)
            cached = (C.c_void_p * max(1, len(a.cached_mains)))(*[c._h for c in a.cached_mains])
            keep += [pv, cached]
            arr[i].air_id = a.air_id
            arr[i].n_cached = len(a.cached_mains)
            arr[i].cached = C.cast(cached, C.POINTER(C.c_void_p))
            if a.cached_traces:
                if len(a.cached_traces) != len(a.cached_mains):
                    raise ValueError("cached_traces must match cached_mains")
                traces = (C.c_void_p * len(a.cached_traces))(*[t._h for t in a.cached_traces])
                keep.append(traces)
                arr[i].cached_traces = C.cast(traces, C.POINTER(C.c_void_p))
            arr[i].common_main = a.common_main._h if a.common_main is not None else None
            arr[i].public_values = pv.ctypes.data_as(C.POINTER(C.c_uint32))
            arr[i].n_public_values = pv.size
        out, ln = C.c_void_p(), C.c_size_t()
        self._check(self._lib.b2s_prove(self._h, pk._h, arr, n, C.byref(out), C.byref(ln)))
        data = C.string_at(out.value, ln.value)
        self._lib.b2s_free(out)
        return data

    def verify(self, pk, proof):
        """StarkEngine::verify(&vk, &proof) (crates/vm/src/arch/vm.rs:1074,1127): returns None, or raises
        VerificationError carrying the reference's error variant name."""
        proof = bytes(proof)
        st = self._lib.b2s_verify(self._h, pk._h, proof, len(proof))
        if st == _lib.ERR_VERIFY_FAILED:
            raise VerificationError((self._lib.b2s_last_error(self._h) or b"").decode())
        self._check(st)

    def proof_hint_stream(self, proof):
        """`proof.write()` of the recursive verifier's Hintable (extensions/native/recursion/src/hints.rs:272-316): the list
        of field-element vectors the next aggregation layer hints to its verifier program (crates/sdk/src/prover/agg.rs:307-321)."""
        words, n = C.POINTER(C.c_uint32)(), C.c_size_t()
        self._check(self._lib.b2s_proof_hint_stream(self._h, bytes(proof), len(proof), C.byref(words), C.byref(n)))
        flat = np.ctypeslib.as_array(words, shape=(n.value,)).copy()
        self._lib.b2s_free(words)
        out, i = [], 1
        for _ in range(int(flat[0])):
            ln = int(flat[i])
            out.append(flat[i + 1:i + 1 + ln].copy())
            i += 1 + ln
        return out

    def proof_witness_stream(self, proof):
        """Witnessable::write of a root-configuration proof (extensions/native/recursion/src/witness.rs:147-190): the witness
        of the static halo2 verifier as (vars [n, 8 limbs], felts [n], exts [n, 4])."""
        words, n = C.POINTER(C.c_uint32)(), C.c_size_t()
        self._check(self._lib.b2s_proof_witness_stream(self._h, bytes(proof), len(proof), C.byref(words), C.byref(n)))
        flat = np.ctypeslib.as_array(words, shape=(n.value,)).copy()
        self._lib.b2s_free(words)
        nv, nf, ne = (int(x) for x in flat[:3])
        a, b = 3 + 8 * nv, 3 + 8 * nv + nf
        return flat[3:a].reshape(nv, 8), flat[a:b], flat[b:b + 4 * ne].reshape(ne, 4)

    def set_pow_witnesses(self, logup_witness=None, fri_witness=None):
        """Replay aid: use the given proof-of-work witnesses instead of searching for the minimum ones."""
        s = 0xFFFFFFFF
        self._check(self._lib.b2s_engine_set_pow_witnesses(self._h, s if logup_witness is None else logup_witness,
                                                           s if fri_witness is None else fri_witness))

    def device(self):
        return self._device

    @property
    def stream(self):
        return int(self._lib.b2s_engine_stream(self._h) or 0)

    def set_stream(self, cuda_stream):
        self._check(self._lib.b2s_engine_set_stream(self._h, cuda_stream))

    def synchronize(self):
        self._check(self._lib.b2s_engine_synchronize(self._h))

    def reset_metrics(self):
        self._check(self._lib.b2s_reset_metrics(self._h))

    def metrics(self):
        m = _lib.Metrics()
        self._check(self._lib.b2s_get_metrics(self._h, C.byref(m)))
        return {k: getattr(m, k) for k, _ in _lib.Metrics._fields_}

    def device_memory(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._check(self._lib.b2s_device_memory(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def bench_modmul(self, threads, iters, ilp=4):
        ms = C.c_double()
        self._check(self._lib.b2s_bench_modmul(self._h, threads, iters, ilp, C.byref(ms)))
        return ms.value

    def bench_poseidon2(self, threads, iters):
        ms = C.c_double()
        self._check(self._lib.b2s_bench_poseidon2(self._h, threads, iters, C.byref(ms)))
        return ms.value

    def _check(self, status):
        if status != 0:
            msg = self._lib.b2s_last_error(self._h)
            raise B2StarkError(status, msg.decode() if msg else "")

    def _track(self, obj):
        self._live.add(obj)

    def close(self):
        """Frees the engine and every matrix / PcsData / key still alive on it (their handles point into the engine)."""
        if self._h:
            for obj in list(self._live):
                obj.free()
            self._live.clear()
            self._lib.b2s_engine_free(self._h)
            self._h = None

    def __del__(self):
        # engines are closed explicitly; matrices may outlive a garbage-collected engine otherwise
        pass
