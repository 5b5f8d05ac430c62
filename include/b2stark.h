/* b2stark — C ABI of the B200-native STARK proving backend (hot path of OpenVM's stark-backend).
 *
 * This header is the drop-in boundary. The reference selects its prover backend at compile time through
 * Rust generics (E: StarkFriEngine, E::PD: ProverDevice = TraceCommitter + QuotientCommitter +
 * OpeningProver + DeviceDataTransporter; /root/reference crates/sdk/src/lib.rs:71-83,127-131;
 * crates/vm/src/arch/vm.rs:334-349). The traits themselves live in the un-vendored dependency
 * openvm-stark-backend @ 002f52e (Cargo.toml:113-117); every entry point below cites the in-tree call
 * site of the trait method it stands in for. A Rust shim forwards those methods here.
 *
 * Conventions
 *   - every function returns int32_t: 0 = OK, negative = b2s_status, positive = raw cudaError_t;
 *     no C++ exception crosses the ABI; b2s_last_error() gives a message.
 *   - field elements are 32-bit words. `repr` says how a buffer crossing the ABI encodes them:
 *       B2S_REPR_CANONICAL  value in [0,p)                       (wire format, codec.rs:309-313)
 *       B2S_REPR_MONTY      value*2^32 mod p, in [0,p)           (Plonky3 BabyBear's in-memory form)
 *     Device memory is always Montgomery form, column-major (element (r,c) at c*height + r), the layout
 *     of the reference's DeviceMatrix (crates/circuits/primitives/cuda/include/primitives/trace_access.h:6-22).
 *   - handles are opaque, library-allocated, released with the matching *_free; an engine is not
 *     thread-safe; distinct engines are independent (one per GPU / per worker thread).
 *   - there is no CPU fallback: every compute entry point runs CUDA kernels on the engine's device.
 */
#ifndef B2STARK_H
#define B2STARK_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2s_engine b2s_engine;
typedef struct b2s_matrix b2s_matrix;     /* device matrix, column-major, Montgomery form */
typedef struct b2s_pcs_data b2s_pcs_data; /* Merkle tree + retained LDE matrices of one commitment */

typedef enum {
    B2S_OK = 0,
    B2S_ERR_BAD_ARG = -1,
    B2S_ERR_SHAPE = -2,
    B2S_ERR_OOM = -3,
    B2S_ERR_NO_DEVICE = -4,
    B2S_ERR_CONSTRAINT_DEGREE = -5,
    B2S_ERR_VERIFY_FAILED = -6,
    B2S_ERR_INTERNAL = -7
} b2s_status;

enum { B2S_LAYOUT_ROW_MAJOR = 0, B2S_LAYOUT_COL_MAJOR = 1 };
enum { B2S_REPR_CANONICAL = 0, B2S_REPR_MONTY = 1 };

/* FriParameters {log_blowup, log_final_poly_len, num_queries, proof_of_work_bits}
 * (openvm_stark_sdk::config::FriParameters; defaults crates/sdk/src/config/mod.rs:15-18,
 *  docs/vocs/docs/pages/specs/security/security-model.mdx:47-53). */
typedef struct {
    uint32_t log_blowup;
    uint32_t log_final_poly_len;
    uint32_t num_queries;
    uint32_t proof_of_work_bits;
} b2s_fri_params;

/* Device milliseconds accumulated since the last b2s_reset_metrics, measured with CUDA events on the engine's stream;
 * *_launches count kernels launched by this library.
 *  - lde / leaf_hash / compress: kernel classes. They run INSIDE commitments, so they break commit_total_ms (and the
 *    *_commit gauges) down; they are not additive with them.
 *  - commit_total_ms: every commitment made through the engine (b2s_commit*, b2s_mmcs_commit, and those inside b2s_prove).
 *  - the seven gauges of the reference (crates/prof/data/metrics.example.json; emitted around
 *    crates/vm/src/arch/vm.rs:1328-1399) for b2s_prove: the first six are DISJOINT intervals of one proof, in the order
 *    they occur, so their sum is <= stark_prove_excluding_trace_time_ms (the remainder: transcript set-up, key lookups,
 *    proof encoding). The FRI-LogUp proof-of-work grind is counted in generate_perm_trace_time_ms. */
typedef struct {
    double lde_ms;              /* iNTT + coset NTT passes                      */
    double leaf_hash_ms;        /* Poseidon2 sponge over LDE rows                */
    double compress_ms;         /* Poseidon2 2-to-1 layers                       */
    double commit_total_ms;
    double main_trace_commit_time_ms;            /* LDE + Merkle tree of the common main traces               */
    double generate_perm_trace_time_ms;          /* LogUp PoW, challenges, after-challenge traces, running sums */
    double perm_trace_commit_time_ms;            /* commitment of the after-challenge traces                   */
    double quotient_poly_compute_time_ms;        /* constraint evaluation on the quotient domain               */
    double quotient_poly_commit_time_ms;         /* commitment of the quotient chunks                          */
    double pcs_opening_time_ms;                  /* values at zeta, reduced openings, FRI commit phase, PoW, queries */
    double ood_eval_ms;                          /* part of pcs_opening: values at zeta (and zeta g), observed by the host */
    double reduced_opening_ms;                   /* part of pcs_opening: reduced openings per LDE height                  */
    double fri_commit_phase_ms;                  /* part of pcs_opening: FRI layers (pack, tree, root, fold)              */
    double stark_prove_excluding_trace_time_ms;  /* whole b2s_prove                                            */
    uint64_t lde_launches;
    uint64_t leaf_hash_launches;
    uint64_t compress_launches;
    uint64_t other_launches;
    uint64_t poseidon2_permutations; /* permutations executed on the device */
    uint64_t quotient_jit_launches;          /* generated straight-line quotient kernels (qjit.cu)        */
    uint64_t quotient_interpreter_launches;  /* quotient_kernel (B2S_QUOTIENT_JIT=0)                      */
} b2s_metrics;

/* ---- engine ------------------------------------------------------------------------------------------
 * E::new(fri_params)            crates/sdk/src/prover/vm/mod.rs:22, crates/vm/src/utils/stark_utils.rs:115 */
int32_t b2s_engine_new(const b2s_fri_params* params, int32_t cuda_device, b2s_engine** out);
void b2s_engine_free(b2s_engine* e);
/* engine.fri_params()           crates/sdk/src/prover/app.rs:139 */
int32_t b2s_engine_fri_params(const b2s_engine* e, b2s_fri_params* out);
/* The hash of the configuration: BabyBearPoseidon2Config (default; digests are 8 BabyBear words) or
 * BabyBearPoseidon2RootConfig (crates/continuations/src/lib.rs:17, crates/sdk/src/prover/root.rs): Poseidon2 over the
 * BN254 scalar field, width 3, leaves hashed with the multi-field sponge of
 * extensions/native/recursion/src/outer_poseidon2.rs:26-63; a digest is ONE scalar, returned as 8 little-endian 32-bit
 * limbs of its canonical value in the same root8 / proof slots. Affects commitments made after the call. */
#define B2S_HASH_POSEIDON2_BABYBEAR 0
#define B2S_HASH_POSEIDON2_BN254 1
int32_t b2s_engine_set_hash(b2s_engine* e, int32_t kind);
int32_t b2s_engine_hash(const b2s_engine* e);
/* message for the last non-zero status returned on this engine (or globally when e == NULL) */
const char* b2s_last_error(const b2s_engine* e);
/* the CUDA stream all work of this engine is ordered on (cudaStream_t as void*); set lets a caller that
 * already owns a stream (e.g. the reference's tracegen) share it. Latency-bound parts of a commit or proof (LDEs of
 * short matrices, row digests of short height groups, quotient / LogUp / reduced-opening kernels) run on internal
 * side streams that fork from and join back into this stream with events: for the caller everything is ordered on
 * this one stream. With several engines per GPU set CUDA_DEVICE_MAX_CONNECTIONS=32 before CUDA initialises. */
void* b2s_engine_stream(const b2s_engine* e);
int32_t b2s_engine_set_stream(b2s_engine* e, void* cuda_stream);
int32_t b2s_engine_synchronize(b2s_engine* e);
int32_t b2s_reset_metrics(b2s_engine* e);
int32_t b2s_get_metrics(b2s_engine* e, b2s_metrics* out); /* synchronises the stream */
/* bytes currently held by this engine's handles / high-water mark */
int32_t b2s_device_memory(const b2s_engine* e, uint64_t* in_use, uint64_t* peak);

/* ---- matrices (DeviceDataTransporter) ----------------------------------------------------------------
 * transport_matrix_to_device(&Arc<RowMajorMatrix>)      crates/vm/src/arch/vm.rs:731-734
 * transport_matrix_from_device_to_host                  crates/vm/src/arch/vm.rs:1307-1312 */
int32_t b2s_matrix_upload(b2s_engine* e, const uint32_t* host, uint64_t height, uint64_t width,
                          int32_t layout, int32_t repr, b2s_matrix** out);
int32_t b2s_matrix_download(b2s_engine* e, const b2s_matrix* m, uint32_t* host, int32_t layout, int32_t repr);
/* zero-copy adoption of a caller-owned column-major Montgomery-form device buffer (the reference's
 * DeviceMatrix{buffer,height,width}; crates/vm/src/system/cuda/program.rs:60,119). Not freed by us. */
int32_t b2s_matrix_adopt_device(b2s_engine* e, void* device_ptr, uint64_t height, uint64_t width, b2s_matrix** out);
/* uninitialised device matrix owned by the library */
int32_t b2s_matrix_alloc(b2s_engine* e, uint64_t height, uint64_t width, b2s_matrix** out);
/* deterministic synthetic contents generated on the device: element(r,c) =
 * splitmix64(seed ^ ((c << 32) | r)) mod p; same generator in oracle/. */
int32_t b2s_matrix_fill_synthetic(b2s_engine* e, b2s_matrix* m, uint64_t seed);
/* deterministic SATISFYING trace of the tests' chain AIR (openvm_b200/airs.py chain_air; width >= 5), generated on the
 * device on the engine's stream: stands in for the reference's device tracegen (crates/vm/src/system/cuda/, out of scope)
 * in bench.py and the tests. pv_out receives the AIR's two public values (canonical). Same values as
 * airs.chain_trace(air, height, seed, hashed=True) on the host. */
int32_t b2s_matrix_fill_chain(b2s_engine* e, b2s_matrix* m, uint64_t seed, uint32_t pv_out[2]);
uint64_t b2s_matrix_height(const b2s_matrix* m);
uint64_t b2s_matrix_width(const b2s_matrix* m);
void* b2s_matrix_device_ptr(const b2s_matrix* m);
void b2s_matrix_free(b2s_matrix* m);

/* ---- commitment (TraceCommitter) ---------------------------------------------------------------------
 * device.commit(&[matrix]) -> (Com, PcsData)            crates/vm/src/arch/vm.rs:735,
 *                                                       crates/vm/src/system/cuda/program.rs:80
 * = TwoAdicFriPcs::commit: per matrix iDFT -> coset DFT on GENERATOR*H_{N*2^b}, rows bit-reversed
 *   (extensions/native/recursion/src/fri/two_adic_pcs.rs:178-186), then one mixed-height Poseidon2
 *   Merkle tree over all LDEs (extensions/native/circuit/src/poseidon2/README.md:9-17).
 * root8 receives the commitment as 8 canonical words. The PcsData retains the LDE matrices. */
int32_t b2s_commit(b2s_engine* e, const b2s_matrix* const* mats, size_t n_mats, uint32_t root8[8],
                   b2s_pcs_data** out);

/* transport_matrix_to_device + commit for one tall host-resident trace, as one pipelined call
 * (crates/vm/src/arch/vm.rs:731-735 issues the two back to back): column groups are uploaded on a second stream
 * while the previous group is extended and hashed, so the PCIe transfer is hidden behind the arithmetic. The result
 * is identical to b2s_matrix_upload + b2s_commit; the trace itself is not kept on the device (the PcsData holds the
 * LDE). `host` should be pinned for the copy to overlap; layout / repr as for b2s_matrix_upload. */
int32_t b2s_commit_host(b2s_engine* e, const uint32_t* host, uint64_t height, uint64_t width, int32_t layout,
                        int32_t repr, uint32_t root8[8], b2s_pcs_data** out);
/* the two halves of commit, exposed for parity tests and for the quotient / FRI commit-phase callers:
 * coset LDE of evaluations over shift_in*H_N (shift_in canonical; 1 for traces) ... */
int32_t b2s_coset_lde(b2s_engine* e, const b2s_matrix* in, uint32_t log_blowup, uint32_t shift_in,
                      b2s_matrix** out);
/* ... and the MMCS tree over already-extended matrices of power-of-two heights (not consumed). */
int32_t b2s_mmcs_commit(b2s_engine* e, const b2s_matrix* const* mats, size_t n_mats, uint32_t root8[8],
                        b2s_pcs_data** out);

/* ---- one commitment across the GPUs of a node (no reference counterpart: TraceCommitter::commit runs on one device,
 * crates/vm/src/arch/vm.rs:735; host side: openvm_b200/sharded.py) ---------------------------
 * Every rank (one process per GPU) holds a COLUMN shard of the trace and, after the exchange, a ROW block of the LDE
 * of all columns, which is one sub-tree of the Merkle tree (rows are stored bit-reversed). The exchange is fused into
 * the LDE: the last NTT pass stores every output row straight into the buffer of the rank that will hash it, over
 * NVLink peer mappings -- no staging buffer, no separate collective.
 *   b2s_shared_alloc   device buffer that other processes can map (cudaMalloc + cudaIpcGetMemHandle, 64-byte handle)
 *   b2s_shared_open    maps a peer's buffer from its handle (cudaIpcOpenMemHandle); b2s_shared_close unmaps it
 *   b2s_coset_lde_scatter  coset LDE of this rank's columns [col0, col0 + in->width) of a trace of `total_width` columns;
 *                      rank g's buffer peer_bufs[g] (this rank's own allocation at its own index) receives LDE rows
 *                      [g R, (g+1) R), R = (height << log_blowup) / n_peers, as a column-major R x total_width matrix.
 * The caller orders the ranks (a barrier after synchronising the stream) before anyone reads its buffer. */
#define B2S_MAX_PEERS 16
int32_t b2s_shared_alloc(b2s_engine* e, uint64_t bytes, void** device_ptr, uint8_t handle64[64]);
int32_t b2s_shared_free(b2s_engine* e, void* device_ptr);
int32_t b2s_shared_open(b2s_engine* e, const uint8_t handle64[64], void** device_ptr);
int32_t b2s_shared_close(b2s_engine* e, void* device_ptr);
int32_t b2s_coset_lde_scatter(b2s_engine* e, const b2s_matrix* in, uint32_t log_blowup, uint32_t shift_in,
                              void* const* peer_bufs, uint32_t n_peers, uint64_t total_width, uint64_t col0);

size_t b2s_pcs_num_matrices(const b2s_pcs_data* d);
uint32_t b2s_pcs_log_max_height(const b2s_pcs_data* d);
/* borrowed view of the i-th retained (LDE) matrix, valid while `d` lives */
const b2s_matrix* b2s_pcs_matrix(const b2s_pcs_data* d, size_t i);
/* Mmcs::open_batch(index): rows of every matrix (caller order, row index >> (log_max - log_h)),
 * concatenated into rows_out (sum of widths words), and log_max_height sibling digests bottom-up into
 * proof_out (8 words each); canonical words. (verifier side: fri/mod.rs:185-299) */
int32_t b2s_pcs_open(b2s_engine* e, const b2s_pcs_data* d, uint64_t index, uint32_t* rows_out,
                     uint32_t* proof_out);
/* digest layer `level` (0 = leaves), n*8 canonical words, node-major; for tests */
int32_t b2s_pcs_layer(b2s_engine* e, const b2s_pcs_data* d, uint32_t level, uint32_t* out);
void b2s_pcs_free(b2s_pcs_data* d);

/* ---- proving key and proof (StarkEngine::prove) ---------------------------------------------------------
 * device.transport_pk_to_device(&pk)                     crates/vm/src/arch/vm.rs:382,
 *                                                        crates/sdk/src/prover/vm/mod.rs:23
 * The reference's MultiStarkProvingKey is a serde type of the un-vendored openvm-stark-backend; the Rust shim
 *  flattens it into this library's key blob: little-endian u32 words
 *   'B2PK', version (2 or 3), pre_hash[8], log_up_pow_bits, num_airs, then per AIR
 *     width_preprocessed, n_cached, cached_widths[], width_common_main, width_after_challenge, log_quotient_degree,
 *     num_public_values, num_exposed_values,
 *     [if width_preprocessed: commit[8], log_height, row-major canonical trace]
 *     num_nodes, nodes[][4] = {kind,a,b,c}, num_constraints, constraint_idx[]
 *     num_interactions, per interaction {bus, type (0 send / 1 receive), count_node, n_fields, field_nodes[]},
 *     num_chunks, per chunk {n, interaction indices[n]}   (FRI-LogUp: one reciprocal-sum column per chunk, then
 *     the running-sum column; width_after_challenge = num_chunks + 1 extension-field columns)
 *   [version 3, after the AIRs: num_height_constraints, per constraint {coefficients[num_airs], threshold <= p} =
 *    vk.trace_height_constraints (extensions/native/recursion/src/types.rs:70): b2s_prove refuses and b2s_verify rejects
 *    heights with sum coefficients[air] * height >= threshold (crates/vm/src/arch/vm.rs:627-652, stark/mod.rs:836-901)]
 * (node kinds: 0 Constant 1 Preprocessed 2 Main 3 Permutation 4 Public 5 Challenge 6 Exposed 7 IsFirstRow
 *  8 IsLastRow 9 IsTransition 10 Add 11 Sub 12 Neg 13 Mul; content of SymbolicExpressionDag as consumed at
 *  extensions/native/recursion/src/stark/mod.rs:1023-1085, types.rs:19-95).
 * Loading uploads and commits the preprocessed traces (the root must equal the key's commit), validates every DAG node,
 * generates each AIR's quotient kernels as straight-line PTX from its constraint DAG (openvm_b200/csrc/qjit.cu; the
 * driver compiles them once per key), and compiles each AIR's interaction list into the program the LogUp trace kernel
 * interprets. With B2S_QUOTIENT_JIT=0 in the environment at load time the constraints run through an interpreter
 * instead (same results; tests/test_gpu_prove.py compares both with the oracle). */
typedef struct b2s_pk b2s_pk;
int32_t b2s_pk_load(b2s_engine* e, const uint32_t* key_words, size_t n_words, b2s_pk** out);
void b2s_pk_free(b2s_pk* pk);

/* AirProvingContext {cached_mains: Vec<CommittedTraceData>, common_main, public_values}
 * (built at crates/vm/src/arch/vm.rs:588-599; GPU instance crates/vm/src/system/cuda/program.rs:121-125).
 * A cached main is passed as the PcsData returned by b2s_commit on that single trace, plus (CommittedTraceData
 * carries both) the trace itself: cached_traces is only read for AIRs with interactions, whose after-challenge
 * trace is generated from the main trace rows, and may be NULL otherwise. */
typedef struct {
    uint32_t air_id;
    uint32_t n_cached;
    const b2s_pcs_data* const* cached;
    const b2s_matrix* const* cached_traces;  /* n_cached device traces (height x cached width), or NULL */
    const b2s_matrix* common_main;      /* NULL when the AIR has no common main */
    const uint32_t* public_values;      /* canonical */
    uint32_t n_public_values;
} b2s_air_ctx;

/* engine.prove(&pk, ctx) -> Proof<SC>                    crates/vm/src/arch/vm.rs:694,1005
 * Transcript order: extensions/native/recursion/src/stark/mod.rs:140-364; opening: fri/two_adic_pcs.rs:41-376,
 * fri/mod.rs:33-170. AIRs must be given in strictly increasing air_id. The proof is returned in the reference's
 * wire format (crates/sdk/src/codec.rs:92-330) in a buffer released with b2s_free. Proof-of-work witnesses are the
 * minimum valid ones (the reference returns any valid witness). AIRs with interactions run the single FRI-LogUp
 * challenge phase (stark/mod.rs:307-358): log_up_pow_bits grinding, two extension challenges, the after-challenge
 * trace of every interacting AIR committed in one tree, exposed cumulative sums in the proof. */
int32_t b2s_prove(b2s_engine* e, const b2s_pk* pk, const b2s_air_ctx* airs, size_t n_airs, uint8_t** proof_bytes,
                  size_t* proof_len);

/* Replay aid: the reference grinds in parallel and returns ANY valid witness, this library the minimum
 * one. To reproduce a reference proof byte for byte, pass the witnesses found in that proof (canonical values);
 * B2S_POW_SEARCH restores the search. A supplied witness that does not satisfy the transcript fails the proof. */
#define B2S_POW_SEARCH 0xFFFFFFFFu
int32_t b2s_engine_set_pow_witnesses(b2s_engine* e, uint32_t logup_witness, uint32_t fri_witness);

/* engine.verify(&vk, &proof) -> Result<(), VerificationError>    crates/vm/src/arch/vm.rs:1074,1127;
 *                                                                crates/sdk/src/lib.rs:735-737
 * Decodes the wire-format proof (crates/sdk/src/codec.rs:332-642) and runs MultiTraceStarkVerifier::verify as
 * restated in-tree (extensions/native/recursion/src/stark/mod.rs:108-1173) with the engine's FRI parameters against
 * the verifying-key part of `pk`. Host-side like the reference's (milliseconds of sequential transcript work).
 * Returns B2S_OK, or B2S_ERR_VERIFY_FAILED with b2s_last_error naming the reference's VerificationError variant
 * (InvalidProofShape, InvalidOpeningArgument, InvalidPowWitness, NonZeroCumulativeSum, OodEvaluationMismatch ...). */
int32_t b2s_verify(b2s_engine* e, const b2s_pk* pk, const uint8_t* proof_bytes, size_t proof_len);
/* The same verdicts with no device and no engine (the verifier is host code): the key blob is parsed on the host, its
 * preprocessed traces skipped and their commitments kept. hash_kind: B2S_HASH_*. */
int32_t b2s_verify_host(const b2s_fri_params* params, int32_t hash_kind, const uint32_t* key_words, size_t n_words,
                        const uint8_t* proof_bytes, size_t proof_len);

/* Proof -> hint stream of the recursive verifier: what the next aggregation layer feeds its verifier program
 * (`proof.write()` of `Hintable for Proof<BabyBearPoseidon2Config>`, extensions/native/recursion/src/hints.rs:272-316,
 * fri/hints.rs:22-163; consumed at crates/sdk/src/prover/agg.rs:307-321). Host-only; `e` may be NULL (then no message
 * is kept for b2s_last_error). Output, released with b2s_free: n_vectors, then per vector its length and its canonical
 * words. Fails with B2S_ERR_VERIFY_FAILED (InvalidProofShape) on bytes that do not decode. */
int32_t b2s_proof_hint_stream(b2s_engine* e, const uint8_t* proof_bytes, size_t proof_len, uint32_t** words,
                              size_t* n_words);
/* A proof made under B2S_HASH_POSEIDON2_BN254 as the witness of the static (halo2) verifier: Witnessable::write for
 * Proof<BabyBearPoseidon2RootConfig> (extensions/native/recursion/src/witness.rs:147-190, fri/witness.rs:61-117), the input
 * of the wrapper circuit (crates/sdk/src/prover/halo2.rs). e may be NULL (host only). *words (b2s_free) holds n_vars,
 * n_felts, n_exts, then the vars (BN254 scalars, 8 little-endian limbs each), the felts, the exts (4 words each). */
int32_t b2s_proof_witness_stream(b2s_engine* e, const uint8_t* proof_bytes, size_t proof_len, uint32_t** words,
                                 size_t* n_words);

/* The quotient kernels of an AIR are generated as PTX from its constraint DAG when the key is loaded and compiled by
 * the driver (openvm_b200/csrc/qjit.cu; B2S_QUOTIENT_JIT=0 selects the interpreter instead). This entry point returns
 * that PTX text for AIR `air_index` of a key blob without touching a device (release with b2s_free): tests assemble
 * it with ptxas for sm_100a. */
int32_t b2s_debug_quotient_ptx(const uint32_t* key_words, size_t n_words, uint32_t log_blowup, uint32_t air_index,
                               char** ptx_out);

/* ---- primitives run on the device, for known-answer / parity tests ------------------------------------ */
/* n states of 16 canonical words, permuted in place on the GPU */
int32_t b2s_poseidon2_permute(b2s_engine* e, uint32_t* host_states, size_t n_states);
/* integer-pipe micro-benchmark: `iters` dependent Montgomery multiplications per thread on
 * `threads` threads; returns device ms (the measured modmul/s is the IMAD "peak" of DESIGN.md) */
/* Poseidon2-BN254 width 3 on n states of 3 x 8 canonical limbs, in place
 * (extensions/native/compiler/src/constraints/halo2/poseidon2_perm.rs:52-82) */
int32_t b2s_poseidon2_bn254_permute(b2s_engine* e, uint32_t* host_states, size_t n_states);
/* the same on the host cores (no device, no engine): the transcript and the verifier of the root configuration use it */
int32_t b2s_host_poseidon2_bn254_permute(uint32_t* states, size_t n_states);
int32_t b2s_bench_modmul(b2s_engine* e, uint64_t threads, uint32_t iters, int32_t ilp, double* ms_out);
/* Poseidon2 permutations/s with state in registers and no memory traffic */
int32_t b2s_bench_poseidon2(b2s_engine* e, uint64_t threads, uint32_t iters, double* ms_out);

/* ---- allocation-free entry point -----------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Poseidon2 Merkle root of ONE already-extended matrix (column-major, Montgomery form, on the CURRENT device,
 * power-of-two height): the same tree b2s_mmcs_commit builds for that single matrix, bit for bit. The digest layers,
 * which b2s_mmcs_commit takes from the engine's pool, live in a file-scope __device__ array of
 * B2S_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit. The call takes no engine, launches on
 * `cuda_stream` (cudaStream_t as void*, NULL = default stream), does not synchronise, and calls no cudaMalloc /
 * cudaFree / allocating routine. d_root8 is DEVICE memory for 8 canonical words. Heights above B2S_STATIC_MAX_ROWS or
 * widths above B2S_STATIC_MAX_WIDTH are refused with B2S_ERR_SHAPE (message: b2s_last_error(NULL)).
 * All calls on a device share the scratch array (it keeps the last call's contents; nothing in it needs clearing,
 * every word read was written earlier in the same call): calls that overlap on the GPU race -- order them on one stream
 * or with events. (openvm_b200/csrc/noalloc.cu) */
#define B2S_STATIC_MAX_ROWS 65536u
#define B2S_STATIC_MAX_WIDTH 4096u
#define B2S_STATIC_SCRATCH_BYTES (4u << 20)
int32_t b2s_merkle_root_static(const uint32_t* d_matrix, uint64_t height, uint64_t width, uint32_t* d_root8,
                               void* cuda_stream);

void b2s_free(void* p); /* releases host buffers returned by the library */

#ifdef __cplusplus
}
#endif
#endif /* B2STARK_H */
