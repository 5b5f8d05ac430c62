// Internal definitions shared by the translation units of libb2stark.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <map>
#include <stdexcept>

#include "../../include/b2stark.h"
#include "field.cuh"

namespace b2 {


struct StatusError {
    int32_t code;
    std::string msg;
};

[[noreturn]] inline void fail(int32_t code, const std::string& msg) { throw StatusError{code, msg}; }

#define B2_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t _e = (expr);                                                                        \
        if (_e != cudaSuccess) {                                                                        \
            ::b2::fail((int32_t)_e, std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                                        ":" + std::to_string(__LINE__) + ")");                          \
        }                                                                                               \
    } while (0)

#define B2_CHECK_LAUNCH() B2_CUDA(cudaGetLastError())

#define B2_REQUIRE(cond, code, msg)                      \
    do {                                                 \
        if (!(cond)) ::b2::fail((code), std::string(msg)); \
    } while (0)

struct TwiddleSet;
struct ScaleSet;

enum Stage { ST_LDE = 0, ST_LEAF, ST_COMPRESS, ST_COMMIT, ST_MAIN_COMMIT, ST_PERM_GEN, ST_PERM_COMMIT, ST_QUOTIENT,
             ST_QUOTIENT_COMMIT, ST_OPENING, ST_OOD, ST_REDUCED, ST_FRI_COMMIT, ST_PROVE, ST_COUNT };

struct StageTimer {
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
    double ms = 0;
};

}  // namespace b2

struct b2s_engine {
    int device = 0;
    int num_sms = 0;                 // cudaDevAttrMultiProcessorCount of `device`, set by b2s_engine_new
    b2s_fri_params params{};
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string last_error;
    uint64_t mem_in_use = 0, mem_peak = 0;
    // Freed device blocks, kept per (stream they were freed on, size) and handed out again for the same stream and size: a
    // proof makes ~1000 allocations of the same sizes every time, and every cudaMallocAsync / cudaFreeAsync is a driver call
    // under the context lock that K engines share. Reuse on the same stream keeps the stream-ordered semantics of
    // cudaFreeAsync (api.cu). Flushed when it exceeds cache_cap, when the stream changes, and with the engine.
    std::map<std::pair<cudaStream_t, size_t>, std::vector<void*>> block_cache;
    size_t cache_bytes = 0, cache_cap = (size_t)48 << 30;
    void flush_cache();
    bool metrics_enabled = true;
    b2::StageTimer timers[b2::ST_COUNT];
    std::vector<cudaEvent_t> event_pool;
    uint64_t launches[4] = {0, 0, 0, 0};  // lde, leaf, compress, other
    uint64_t quotient_launches[2] = {0, 0};  // generated kernels, interpreter
    uint64_t permutations = 0;
    cudaStream_t copy_stream = nullptr;                     // host -> device staging of the streaming commit
    cudaStream_t hash_stream = nullptr;                     // row digests of short height groups next to the tall levels (merkle.cu)
    std::vector<cudaStream_t> aux_streams;                  // side streams for the quotient kernels of a proof (qjit.cu)
    std::map<uint64_t, uint32_t*> selector_cache;           // (log height, log quotient degree, blowup) -> 4 planes
    size_t selector_cache_bytes = 0;
    uint32_t pow_override[2] = {0xFFFFFFFFu, 0xFFFFFFFFu};  // [logup, fri]; 0xFFFFFFFF = search for the minimum witness
    std::map<unsigned, b2::TwiddleSet*> twiddles;  // by log size
    std::map<uint64_t, b2::ScaleSet*> scales;      // by (logn, log_blowup, shift_in)
    unsigned* tile_ctr = nullptr;                  // work counter of the persistent leaf-hash kernel (merkle.cu)
    int hash_kind = 0;                             // B2S_HASH_*: 0 Poseidon2-BabyBear-16 digests of 8 words, 1 Poseidon2-BN254-3 (root config)

    void* pinned = nullptr;                        // grow-only pinned host buffer for small asynchronous read-backs
    size_t pinned_bytes = 0;
    void* host_pinned(size_t bytes);
    void* dmalloc(size_t bytes);
    void dfree(void* p, size_t bytes);
    cudaEvent_t get_event();
    void stage_begin(b2::Stage s);
    void stage_end(b2::Stage s);
    void collect_timers();
};

struct b2s_matrix {
    b2s_engine* eng = nullptr;
    uint32_t* d = nullptr;  // column-major, Montgomery form
    uint64_t height = 0, width = 0;
    bool owned = true;
    size_t bytes() const { return (size_t)height * width * 4; }
};

struct b2s_pcs_data {
    b2s_engine* eng = nullptr;
    std::vector<b2s_matrix*> mats;      // caller order; LDE matrices
    std::vector<bool> owns;             // whether mats[i] is freed with the tree
    unsigned log_max_height = 0;
    // digest layers, level 0 = leaves; each layer is 8 word-planes of n nodes: word w of node i at [w*n + i]
    std::vector<uint32_t*> layers;
    std::vector<uint64_t> layer_nodes;
    std::vector<const uint32_t*> col_ptrs;  // host copy of the column pointer table (source of an asynchronous upload)
};

namespace b2 {
// one AIR of a loaded proving key (prover.cu)
struct AirKeyDev {
    uint32_t width_prep = 0, width_common = 0, width_after = 0, log_qd = 0, num_pv = 0, num_exposed = 0;
    std::vector<uint32_t> cached_widths;
    uint32_t prep_log_height = 0;
    uint32_t prep_commit[8] = {0};       // Montgomery
    b2s_matrix* prep_trace = nullptr;
    b2s_pcs_data* prep_pcs = nullptr;
    // compiled constraint program: {op, dst, a, b} words, device copy + host count
    std::vector<uint4> program;
    uint4* d_program = nullptr;
    uint32_t num_slots = 0, num_slots_e = 0, num_constraints = 0;
    // FRI-LogUp: interaction program (evaluates message fields / counts on trace rows), per-interaction bus and
    // message length, number of chunks (= after-challenge EF columns - 1)
    std::vector<uint4> iprogram;
    uint4* d_iprogram = nullptr;
    uint32_t islots = 0, num_chunks = 0, max_fields = 0;
    std::vector<uint32_t> inter_bus, inter_len;
    // symbolic constraints as loaded ({kind, a, b, c} per node), kept for the verifier's evaluation at zeta
    std::vector<uint4> nodes;
    std::vector<uint32_t> constraint_idx;
    // JIT-compiled quotient kernels (qjit.cu): cudaKernel_t handles, empty when the interpreter is used
    std::vector<void*> jit_kernels;
    std::vector<std::string> jit_ptx;     // only while the key is being loaded
    uint32_t jit_num_mats = 0;
    uint32_t total_cols() const {
        uint32_t t = width_prep + width_common;
        for (uint32_t w : cached_widths) t += w;
        return t;
    }
};
}  // namespace b2

struct b2s_pk {
    b2s_engine* eng = nullptr;
    uint32_t pre_hash[8] = {0};          // layer form: Montgomery words, or (hash_kind 1) canonical limbs of one scalar
    int hash_kind = 0;                   // the engine's hash when the key was loaded (preprocessed commitments depend on it)
    uint32_t log_up_pow_bits = 0;
    std::vector<b2::AirKeyDev> airs;
    // LinearConstraint list of the verifying key (extensions/native/recursion/src/types.rs:70): for every constraint,
    // sum_i coefficients[air i] * height_i < threshold over the AIRs present in a proof (stark/mod.rs:836-901, vm.rs:627-652)
    std::vector<std::vector<uint32_t>> height_coefficients;
    std::vector<uint32_t> height_thresholds;
    // index of the first violated constraint, or -1
    template <class Ids, class Logs>
    int violated_height_constraint(const Ids& air_ids, const Logs& log_heights, size_t n) const {
        for (size_t c = 0; c < height_thresholds.size(); c++) {
            unsigned __int128 sum = 0;
            for (size_t k = 0; k < n; k++) sum += (unsigned __int128)height_coefficients[c][air_ids[k]] << log_heights[k];
            if (sum >= height_thresholds[c]) return (int)c;
        }
        return -1;
    }
};

namespace b2 {

inline unsigned log2_exact(uint64_t x) {
    unsigned l = 0;
    while (((uint64_t)1 << l) < x) l++;
    return l;
}
inline bool is_pow2(uint64_t x) { return x && !(x & (x - 1)); }

// RAII stage scope
struct StageScope {
    b2s_engine* e;
    Stage s;
    StageScope(b2s_engine* e_, Stage s_) : e(e_), s(s_) { e->stage_begin(s); }
    ~StageScope() { e->stage_end(s); }
};

// ---- kernels' host entry points (defined in the .cu files) ---------------------------------------------
// ntt.cu
#define B2_MAX_PEERS 16
// row-sharded LDE output over peer memory (sharded commitment): see PassParams in ntt.cu
struct PeerScatter {
    uint32_t* peer[B2_MAX_PEERS];   // per rank: column-major (2^log_rows) x total_width buffer, mapped into this process
    uint32_t n_peers;
    uint32_t log_rows;               // rows per rank
    uint64_t col0;                   // global index of the first column of this shard
};
void lde_columns(b2s_engine* e, const uint32_t* in, uint64_t in_col_stride, unsigned logn, uint64_t width,
                 unsigned log_blowup, Fp shift_in, uint32_t* out, uint64_t out_col_stride, const PeerScatter* ps = nullptr);
void free_twiddles(b2s_engine* e);
void free_scales(b2s_engine* e);
// merkle.cu
void mmcs_build(b2s_engine* e, b2s_pcs_data* d, uint32_t* leaf_layer = nullptr);
// merkle_bn254.cu: the root configuration's hash (one BN254 scalar per digest, canonical limbs in the same 8 planes)
void bn_launch_level(b2s_engine* e, const uint32_t* const* cols, uint32_t ncols, uint64_t n, const uint32_t* prev, uint32_t* out,
                     int mode);
void bn_leaf_absorb(b2s_engine* e, const uint32_t* base, uint32_t ncols, uint64_t n, uint32_t* state, uint32_t* out, bool first,
                    bool last);
void bn_permute_states(b2s_engine* e, uint32_t* d_states, uint64_t n);
// proof-of-work search over the multi-field transcript: state = 3 x 8 Montgomery limbs, pending = canonical observed words
void bn_pow_grind(b2s_engine* e, const uint32_t state24[24], const uint32_t pending[6], uint32_t n_pending, uint32_t bits,
                  uint32_t base, uint32_t count, uint32_t* d_result);

void leaf_absorb(b2s_engine* e, const uint32_t* base, uint32_t ncols, uint64_t n, uint32_t* state, uint32_t* out, bool first,
                 bool last);
void poseidon2_permute_states(b2s_engine* e, uint32_t* d_states, size_t n);
void convert_repr(b2s_engine* e, uint32_t* d, size_t n, bool to_monty);
void transpose_to_colmajor(b2s_engine* e, const uint32_t* d_rowmajor, uint32_t* d_colmajor, uint64_t height,
                           uint64_t width, bool to_monty);
void transpose_to_rowmajor(b2s_engine* e, const uint32_t* d_colmajor, uint32_t* d_rowmajor, uint64_t height,
                           uint64_t width, bool to_canonical);
void fill_synthetic(b2s_engine* e, uint32_t* d, uint64_t height, uint64_t width, uint64_t seed);
void fill_chain(b2s_engine* e, uint32_t* d, uint64_t height, uint64_t width, uint64_t seed, uint32_t pv[2]);
void gather_open(b2s_engine* e, const b2s_pcs_data* d, uint64_t index, uint32_t* h_rows, uint32_t* h_proof);
void download_layer(b2s_engine* e, const b2s_pcs_data* d, uint32_t level, uint32_t* h_out);
double bench_modmul(b2s_engine* e, uint64_t threads, uint32_t iters, int ilp);
struct OpenDesc {
    const uint32_t* base;  // column 0 of a matrix, or plane 0 of a digest layer
    uint64_t stride;       // height (matrix) or node count (layer)
    uint64_t row;
    uint32_t count;        // width or 8; OPEN_RAW set: words are copied as they are (BN254 digest limbs), not Montgomery -> canonical
    uint32_t out_off;
};
constexpr uint32_t OPEN_RAW = 0x80000000u;
inline uint32_t digest_words_desc(const b2s_engine* e) { return e->hash_kind ? (8u | OPEN_RAW) : 8u; }
void gather_rows(b2s_engine* e, const std::vector<OpenDesc>& descs, size_t total_words, uint32_t* h_out);
// qjit.cu
bool quotient_jit_enabled();
size_t quotient_jit_prepare(AirKeyDev& a);
void quotient_jit_compile(AirKeyDev& a, size_t k);
std::string quotient_jit_ptx(const AirKeyDev& a);
// all quotient kernels of one proof between one fork and one join over the engine's side streams (qjit.cu)
struct QuotientBatch {
    struct Finish { const uint32_t* acc; uint32_t K; const uint32_t* sel; uint32_t* out; unsigned n, lq; };
    b2s_engine* e = nullptr;
    unsigned next = 0;
    std::vector<Finish> finishes;
};
void ensure_aux_streams(b2s_engine* e);
cudaStream_t aux_fork(b2s_engine* e, unsigned k);   // side stream k, ordered behind the engine's stream as of now
void aux_join(b2s_engine* e);                       // the engine's stream waits for every side stream
void quotient_batch_begin(b2s_engine* e, QuotientBatch& qb);
void quotient_jit_launch(QuotientBatch& qb, const AirKeyDev& a, const uint32_t* const* mats, const uint32_t* publics,
                         const uint32_t* alpha_pows, const uint32_t* ef_consts, uint32_t* acc, uint32_t* out, unsigned n, unsigned lq);
void quotient_batch_end(QuotientBatch& qb);
void free_selectors(b2s_engine* e);
// prover.cu
b2s_pk* pk_load(b2s_engine* e, const uint32_t* words, size_t n, bool host_only = false, uint32_t log_blowup = 0);
void pk_free(b2s_pk* pk);
void verify(const b2s_engine* e, const b2s_pk* pk, const uint8_t* proof, size_t len);
std::vector<uint32_t> proof_hint_stream(const uint8_t* proof, size_t len);
std::vector<uint32_t> proof_witness_stream(const uint8_t* bytes, size_t len);
void prove(b2s_engine* e, const b2s_pk* pk, const b2s_air_ctx* airs, size_t n_airs, std::vector<uint8_t>& out);
// api.cu helpers shared with the prover
b2s_matrix* new_matrix(b2s_engine* e, uint64_t height, uint64_t width);
void free_pcs(b2s_pcs_data* d);
// commit `mats` (evaluations over shifts[i]*H, natural order): LDE each, then one MMCS tree; the PcsData owns the LDEs
b2s_pcs_data* commit_matrices(b2s_engine* e, const b2s_matrix* const* mats, const Fp* shifts, size_t n);
double bench_poseidon2(b2s_engine* e, uint64_t threads, uint32_t iters);

}  // namespace b2
