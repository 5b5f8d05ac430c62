"""ctypes binding of include/b2stark.h. Fails loudly when the CUDA library is missing (no CPU fallback)."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libb2stark.so")

u32p = C.POINTER(C.c_uint32)
vp = C.c_void_p


class FriParams(C.Structure):
    _fields_ = [("log_blowup", C.c_uint32), ("log_final_poly_len", C.c_uint32), ("num_queries", C.c_uint32),
                ("proof_of_work_bits", C.c_uint32)]


class Metrics(C.Structure):
    _fields_ = [("lde_ms", C.c_double), ("leaf_hash_ms", C.c_double), ("compress_ms", C.c_double),
                ("commit_total_ms", C.c_double), ("main_trace_commit_time_ms", C.c_double),
                ("generate_perm_trace_time_ms", C.c_double), ("perm_trace_commit_time_ms", C.c_double),
                ("quotient_poly_compute_time_ms", C.c_double), ("quotient_poly_commit_time_ms", C.c_double),
                ("pcs_opening_time_ms", C.c_double), ("ood_eval_ms", C.c_double), ("reduced_opening_ms", C.c_double),
                ("fri_commit_phase_ms", C.c_double), ("stark_prove_excluding_trace_time_ms", C.c_double),
                ("lde_launches", C.c_uint64), ("leaf_hash_launches", C.c_uint64),
                ("compress_launches", C.c_uint64), ("other_launches", C.c_uint64),
                ("poseidon2_permutations", C.c_uint64), ("quotient_jit_launches", C.c_uint64),
                ("quotient_interpreter_launches", C.c_uint64)]


ERR_VERIFY_FAILED = -6


class AirCtx(C.Structure):
    _fields_ = [("air_id", C.c_uint32), ("n_cached", C.c_uint32), ("cached", C.POINTER(C.c_void_p)),
                ("cached_traces", C.POINTER(C.c_void_p)), ("common_main", C.c_void_p), ("public_values", C.POINTER(C.c_uint32)),
                ("n_public_values", C.c_uint32)]


# name -> (restype, argtypes); must list every symbol declared in include/b2stark.h
SIGNATURES = {
    "b2s_engine_new": (C.c_int32, [C.POINTER(FriParams), C.c_int32, C.POINTER(vp)]),
    "b2s_engine_free": (None, [vp]),
    "b2s_engine_fri_params": (C.c_int32, [vp, C.POINTER(FriParams)]),
    "b2s_last_error": (C.c_char_p, [vp]),
    "b2s_engine_stream": (vp, [vp]),
    "b2s_engine_set_stream": (C.c_int32, [vp, vp]),
    "b2s_engine_synchronize": (C.c_int32, [vp]),
    "b2s_reset_metrics": (C.c_int32, [vp]),
    "b2s_get_metrics": (C.c_int32, [vp, C.POINTER(Metrics)]),
    "b2s_device_memory": (C.c_int32, [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "b2s_matrix_upload": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, C.POINTER(vp)]),
    "b2s_matrix_download": (C.c_int32, [vp, vp, vp, C.c_int32, C.c_int32]),
    "b2s_matrix_adopt_device": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint64, C.POINTER(vp)]),
    "b2s_matrix_alloc": (C.c_int32, [vp, C.c_uint64, C.c_uint64, C.POINTER(vp)]),
    "b2s_matrix_fill_synthetic": (C.c_int32, [vp, vp, C.c_uint64]),
    "b2s_matrix_fill_chain": (C.c_int32, [vp, vp, C.c_uint64, u32p]),
    "b2s_matrix_height": (C.c_uint64, [vp]),
    "b2s_matrix_width": (C.c_uint64, [vp]),
    "b2s_matrix_device_ptr": (vp, [vp]),
    "b2s_matrix_free": (None, [vp]),
    "b2s_commit": (C.c_int32, [vp, C.POINTER(vp), C.c_size_t, u32p, C.POINTER(vp)]),
    "b2s_coset_lde": (C.c_int32, [vp, vp, C.c_uint32, C.c_uint32, C.POINTER(vp)]),
    "b2s_mmcs_commit": (C.c_int32, [vp, C.POINTER(vp), C.c_size_t, u32p, C.POINTER(vp)]),
    "b2s_proof_hint_stream": (C.c_int32, [vp, C.c_char_p, C.c_size_t, C.POINTER(C.POINTER(C.c_uint32)), C.POINTER(C.c_size_t)]),
    "b2s_proof_witness_stream": (C.c_int32, [vp, C.c_char_p, C.c_size_t, C.POINTER(C.POINTER(C.c_uint32)), C.POINTER(C.c_size_t)]),
    "b2s_shared_alloc": (C.c_int32, [vp, C.c_uint64, C.POINTER(vp), C.c_char_p]),
    "b2s_shared_free": (C.c_int32, [vp, vp]),
    "b2s_shared_open": (C.c_int32, [vp, C.c_char_p, C.POINTER(vp)]),
    "b2s_shared_close": (C.c_int32, [vp, vp]),
    "b2s_coset_lde_scatter": (C.c_int32, [vp, vp, C.c_uint32, C.c_uint32, C.POINTER(vp), C.c_uint32, C.c_uint64, C.c_uint64]),
    "b2s_pcs_num_matrices": (C.c_size_t, [vp]),
    "b2s_pcs_log_max_height": (C.c_uint32, [vp]),
    "b2s_pcs_matrix": (vp, [vp, C.c_size_t]),
    "b2s_pcs_open": (C.c_int32, [vp, vp, C.c_uint64, vp, vp]),
    "b2s_pcs_layer": (C.c_int32, [vp, vp, C.c_uint32, vp]),
    "b2s_pcs_free": (None, [vp]),
    "b2s_pk_load": (C.c_int32, [vp, vp, C.c_size_t, C.POINTER(vp)]),
    "b2s_pk_free": (None, [vp]),
    "b2s_commit_host": (C.c_int32, [vp, vp, C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, vp, C.POINTER(vp)]),
    "b2s_engine_set_pow_witnesses": (C.c_int32, [vp, C.c_uint32, C.c_uint32]),
    "b2s_debug_quotient_ptx": (C.c_int32, [vp, C.c_size_t, C.c_uint32, C.c_uint32, C.POINTER(vp)]),
    "b2s_verify": (C.c_int32, [vp, vp, C.c_char_p, C.c_size_t]),
    "b2s_prove": (C.c_int32, [vp, vp, vp, C.c_size_t, C.POINTER(vp), C.POINTER(C.c_size_t)]),
    "b2s_poseidon2_permute": (C.c_int32, [vp, vp, C.c_size_t]),
    "b2s_poseidon2_bn254_permute": (C.c_int32, [vp, vp, C.c_size_t]),
    "b2s_verify_host": (C.c_int32, [C.c_void_p, C.c_int32, vp, C.c_size_t, C.c_char_p, C.c_size_t]),
    "b2s_host_poseidon2_bn254_permute": (C.c_int32, [vp, C.c_size_t]),
    "b2s_engine_set_hash": (C.c_int32, [vp, C.c_int32]),
    "b2s_engine_hash": (C.c_int32, [vp]),
    "b2s_bench_modmul": (C.c_int32, [vp, C.c_uint64, C.c_uint32, C.c_int32, C.POINTER(C.c_double)]),
    "b2s_bench_poseidon2": (C.c_int32, [vp, C.c_uint64, C.c_uint32, C.POINTER(C.c_double)]),
    "b2s_merkle_root_static": (C.c_int32, [vp, C.c_uint64, C.c_uint64, vp, vp]),
    "b2s_free": (None, [vp]),
}

_lib = None


def load():
    """Load libb2stark.so and bind every declared symbol. Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m openvm_b200.build` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
