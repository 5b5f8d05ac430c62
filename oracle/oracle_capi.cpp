// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header). C ABI for ctypes (tests/, bench cpu_baseline).
#include "babybear.hpp"
#include "poseidon2.hpp"
#include "mmcs.hpp"
#include "dft.hpp"
#include <cstring>

using namespace orc;

static inline EF load_ef(const uint32_t* a) { return EF{F(a[0]), F(a[1]), F(a[2]), F(a[3])}; }

extern "C" {

// ---- constants -------------------------------------------------------------------------------
// out: 4*16 ext_init, 13 partial, 4*16 ext_term, 16 diag  (canonical u32) = 157 words
void orc_poseidon2_constants(uint32_t* out) {
    const Poseidon2Constants& c = poseidon2_constants();
    size_t k = 0;
    for (int r = 0; r < 4; r++) for (int i = 0; i < 16; i++) out[k++] = c.ext_init[r][i].v;
    for (int r = 0; r < 13; r++) out[k++] = c.partial[r].v;
    for (int r = 0; r < 4; r++) for (int i = 0; i < 16; i++) out[k++] = c.ext_term[r][i].v;
    for (int i = 0; i < 16; i++) out[k++] = c.diag[i].v;
}
uint32_t orc_two_adic_generator(uint32_t bits) { return two_adic_generator(bits).v; }

// ---- field ------------------------------------------------------------------------------------
void orc_ef_mul(const uint32_t* a, const uint32_t* b, uint32_t* out) {
    EF x = load_ef(a), y = load_ef(b);
    EF r = x * y;
    for (int i = 0; i < 4; i++) out[i] = r.c[i].v;
}
void orc_ef_inv(const uint32_t* a, uint32_t* out) {
    EF x = load_ef(a);
    EF r = einv(x);
    for (int i = 0; i < 4; i++) out[i] = r.c[i].v;
}

// ---- poseidon2 --------------------------------------------------------------------------------
void orc_poseidon2_permute(uint32_t* state, size_t n_states) {
    for (size_t k = 0; k < n_states; k++) {
        F s[16];
        for (int i = 0; i < 16; i++) s[i] = F(state[16 * k + i]);
        poseidon2_permute(s);
        for (int i = 0; i < 16; i++) state[16 * k + i] = s[i].v;
    }
}
void orc_hash_slice(const uint32_t* x, size_t n, uint32_t* out8) {
    Digest d = hash_slice(reinterpret_cast<const F*>(x), n);
    for (int i = 0; i < 8; i++) out8[i] = d[i].v;
}
void orc_compress(const uint32_t* l, const uint32_t* r, uint32_t* out8) {
    Digest a, b;
    for (int i = 0; i < 8; i++) { a[i] = F(l[i]); b[i] = F(r[i]); }
    Digest d = compress(a, b);
    for (int i = 0; i < 8; i++) out8[i] = d[i].v;
}

// ---- dft / lde --------------------------------------------------------------------------------
void orc_dft(uint32_t* a, uint32_t logn, int inverse) {
    if (inverse) idft(reinterpret_cast<F*>(a), logn); else dft(reinterpret_cast<F*>(a), logn);
}
// column-major in (height 2^logn) -> column-major out (height 2^(logn+b)), bit-reversed rows
void orc_coset_lde(const uint32_t* in, uint32_t logn, uint32_t width, uint32_t log_blowup, uint32_t shift_in,
                   uint32_t* out, int threads) {
    size_t n = (size_t)1 << logn, m = n << log_blowup;
    parallel_for(width, threads, [&](size_t lo, size_t hi) {
        for (size_t c = lo; c < hi; c++)
            coset_lde_column(reinterpret_cast<const F*>(in) + c * n, logn, log_blowup, F(shift_in),
                             reinterpret_cast<F*>(out) + c * m);
    });
}
uint32_t orc_eval_naive(const uint32_t* evals, uint32_t logn, uint32_t shift_in, uint32_t x) {
    return eval_poly_from_evals_naive(reinterpret_cast<const F*>(evals), logn, F(shift_in), F(x)).v;
}

// ---- mmcs -------------------------------------------------------------------------------------
// n matrices (column-major, canonical u32). Returns root; if layers_out != NULL it receives all digest
// layers concatenated bottom-up (hmax + hmax/2 + ... + 1 digests of 8 words).
void orc_mmcs_commit(const uint32_t* const* mats, const uint64_t* heights, const uint64_t* widths, size_t n,
                     uint32_t* root8, uint32_t* layers_out, int threads) {
    std::vector<MatView> ms(n);
    for (size_t i = 0; i < n; i++) ms[i] = MatView{reinterpret_cast<const F*>(mats[i]), (size_t)heights[i], (size_t)widths[i]};
    MerkleTree t;
    t.build(ms, threads);
    Digest r = t.root();
    for (int i = 0; i < 8; i++) root8[i] = r[i].v;
    if (layers_out) {
        size_t k = 0;
        for (auto& L : t.layers) for (auto& d : L) for (int i = 0; i < 8; i++) layers_out[k++] = d[i].v;
    }
}

// open + verify round trip at `index`; returns 1 if the oracle verifier accepts the oracle opening
int orc_mmcs_open_verify(const uint32_t* const* mats, const uint64_t* heights, const uint64_t* widths, size_t n,
                         uint64_t index, uint32_t* proof_out /* log_max_height*8 */, int corrupt) {
    std::vector<MatView> ms(n);
    std::vector<size_t> hs(n);
    for (size_t i = 0; i < n; i++) { ms[i] = MatView{reinterpret_cast<const F*>(mats[i]), (size_t)heights[i], (size_t)widths[i]}; hs[i] = heights[i]; }
    MerkleTree t;
    t.build(ms, 1);
    std::vector<std::vector<F>> rows;
    std::vector<Digest> proof;
    t.open(index, rows, proof);
    if (proof_out) { size_t k = 0; for (auto& d : proof) for (int i = 0; i < 8; i++) proof_out[k++] = d[i].v; }
    if (corrupt && !rows.empty() && !rows[0].empty()) rows[0][0] += F(1);
    return mmcs_verify_batch(t.root(), hs, index, rows, proof) ? 1 : 0;
}

// Full commit of one or more trace matrices (all on the trivial coset, shift 1):
// LDE each, then MMCS over the LDEs. lde_out[i] may be NULL.
void orc_commit(const uint32_t* const* mats, const uint64_t* log_heights, const uint64_t* widths, size_t n,
                uint32_t log_blowup, uint32_t* root8, uint32_t* const* lde_out, int threads) {
    std::vector<std::vector<F>> ldes(n);
    std::vector<MatView> ms(n);
    for (size_t i = 0; i < n; i++) {
        size_t h = (size_t)1 << log_heights[i], m = h << log_blowup;
        ldes[i].resize(m * widths[i]);
        orc_coset_lde(mats[i], (uint32_t)log_heights[i], (uint32_t)widths[i], log_blowup, 1,
                      reinterpret_cast<uint32_t*>(ldes[i].data()), threads);
        ms[i] = MatView{ldes[i].data(), m, (size_t)widths[i]};
        if (lde_out && lde_out[i]) memcpy(lde_out[i], ldes[i].data(), m * widths[i] * 4);
    }
    MerkleTree t;
    t.build(ms, threads);
    Digest r = t.root();
    for (int i = 0; i < 8; i++) root8[i] = r[i].v;
}

// ---- synthetic input (same generator as b2s_matrix_fill_synthetic) ---------
static inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// column-major canonical output
void orc_fill_synthetic(uint32_t* out, uint64_t height, uint64_t width, uint64_t seed) {
    for (uint64_t c = 0; c < width; c++)
        for (uint64_t r = 0; r < height; r++) out[c * height + r] = (uint32_t)(splitmix64(seed ^ ((c << 32) | r)) % P);
}

}  // extern "C"

// ---- STARK prove / verify (oracle/stark.hpp, oracle/stark_verify.hpp) -------------------------------------
#include "stark_verify.hpp"
#include "hints.hpp"
#include "witness.hpp"
static std::string g_orc_error;
extern "C" {
const char* orc_last_error() { return g_orc_error.c_str(); }
void orc_free(void* p) { free(p); }

// mats: for every AIR in order, its cached main partitions then (if the key gives it a common main) the common
// main; column-major canonical words, `heights[i]` rows for AIR i. Returns 0 and a malloc'ed proof on success.
int orc_stark_prove(const uint32_t* key_words, size_t key_n, size_t n_airs, const uint32_t* air_ids,
                    const uint32_t* const* mats, const uint64_t* heights, const uint32_t* const* pvs,
                    uint32_t log_blowup, uint32_t num_queries, uint32_t pow_bits, int threads, uint8_t** proof_out,
                    size_t* len_out) {
    try {
        KeyBlob key = parse_key(key_words, key_n);
        std::vector<AirInput> airs(n_airs);
        size_t mi = 0;
        for (size_t i = 0; i < n_airs; i++) {
            AirInput& a = airs[i];
            a.air_id = air_ids[i];
            if (a.air_id >= key.airs.size()) throw std::runtime_error("air_id out of range");
            const AirKey& k = key.airs[a.air_id];
            auto load = [&](uint32_t w) {
                ColMatrix m(heights[i], w);
                const F* src = reinterpret_cast<const F*>(mats[mi++]);
                std::copy(src, src + m.d.size(), m.d.begin());
                return m;
            };
            for (uint32_t w : k.cached_widths) a.cached.push_back(load(w));
            if (k.width_common) a.common = load(k.width_common);
            for (uint32_t j = 0; j < k.num_pv; j++) a.public_values.push_back(F(pvs[i][j]));
        }
        Proof p = stark_prove(key, airs, FriParams{log_blowup, num_queries, pow_bits}, threads);
        std::vector<uint8_t> bytes = encode_proof(p);
        *proof_out = (uint8_t*)malloc(bytes.size());
        memcpy(*proof_out, bytes.data(), bytes.size());
        *len_out = bytes.size();
        return 0;
    } catch (const std::exception& ex) {
        g_orc_error = ex.what();
        return -1;
    }
}

// 0 = accepted; -1 = rejected / malformed (orc_last_error names the reference error it corresponds to)
int orc_stark_verify(const uint32_t* key_words, size_t key_n, const uint8_t* proof, size_t len, uint32_t log_blowup,
                     uint32_t num_queries, uint32_t pow_bits) {
    try {
        KeyBlob key = parse_key(key_words, key_n);
        Proof p = decode_proof(proof, len);
        // decode -> encode must round-trip (crates/sdk/tests/integration_test.rs:405-425)
        std::vector<uint8_t> again = encode_proof(p);
        if (again.size() != len || memcmp(again.data(), proof, len) != 0) throw VerifyError("codec round trip differs");
        stark_verify(key, p, FriParams{log_blowup, num_queries, pow_bits});
        return 0;
    } catch (const std::exception& ex) {
        g_orc_error = ex.what();
        return -1;
    }
}

// challenger known-answer hook: observe `n_obs` values, then sample `n_samples` (tests drive this against the
// device/host challenger of the product library)
void orc_challenger_run(const uint32_t* obs, size_t n_obs, uint32_t* samples, size_t n_samples) {
    Challenger ch;
    for (size_t i = 0; i < n_obs; i++) ch.observe(F(obs[i]));
    for (size_t i = 0; i < n_samples; i++) samples[i] = ch.sample().v;
}
// proof bytes -> hint stream (oracle/hints.hpp), flattened as n_vectors, then per vector its length and words
int orc_proof_hint_stream(const uint8_t* proof, size_t len, uint32_t** words_out, size_t* n_out) {
    try {
        const Stream st = hint_stream(decode_proof(proof, len));
        std::vector<uint32_t> flat;
        flat.push_back((uint32_t)st.size());
        for (auto& v : st) {
            flat.push_back((uint32_t)v.size());
            flat.insert(flat.end(), v.begin(), v.end());
        }
        *words_out = (uint32_t*)malloc(flat.size() * 4 + 4);
        memcpy(*words_out, flat.data(), flat.size() * 4);
        *n_out = flat.size();
        return 0;
    } catch (const std::exception& ex) {
        g_orc_error = ex.what();
        return -1;
    }
}

// configuration switch: 0 BabyBearPoseidon2Config, 1 BabyBearPoseidon2RootConfig (oracle/bn254.hpp); every hash, tree,
// transcript, prover and verifier entry above follows it
void orc_set_hash(int kind) { hash_kind() = kind; }
int orc_get_hash() { return hash_kind(); }
// Poseidon2-BN254 width 3 on n states of 3 x 8 canonical 32-bit limbs, in place
void orc_bn254_permute(uint32_t* st, size_t n) {
    for (size_t i = 0; i < n; i++) {
        Bn s[3];
        for (int k = 0; k < 3; k++) {
            uint64_t w[4];
            for (int j = 0; j < 4; j++) w[j] = (uint64_t)st[24 * i + 8 * k + 2 * j] | ((uint64_t)st[24 * i + 8 * k + 2 * j + 1] << 32);
            s[k] = bn::from_words(w);
        }
        bn::permute(s);
        for (int k = 0; k < 3; k++) {
            uint64_t w[4];
            bn::to_words(s[k], w);
            for (int j = 0; j < 8; j++) st[24 * i + 8 * k + j] = (uint32_t)(w[j / 2] >> (32 * (j & 1)));
        }
    }
}
// transcript hook with digests: ops[i] = 0 observe value a[i], 1 observe the digest at a[i] (index into digests, 8 words
// each), 2 sample into out
void orc_challenger_script(const uint32_t* ops, const uint32_t* args, size_t n_ops, const uint32_t* digests, uint32_t* out) {
    Challenger ch;
    size_t k = 0;
    for (size_t i = 0; i < n_ops; i++) {
        if (ops[i] == 0) ch.observe(F(args[i]));
        else if (ops[i] == 1) {
            Digest d;
            for (int w = 0; w < 8; w++) d[w].v = digests[8 * args[i] + w];
            ch.observe_digest(d);
        } else out[k++] = ch.sample().v;
    }
}

// root proof bytes -> witness streams (oracle/witness.hpp), flattened as n_vars, n_felts, n_exts, then the vars (8 limbs
// each), the felts, the exts (4 words each). Decoded under the root configuration whatever the switch says.
int orc_proof_witness_stream(const uint8_t* proof, size_t len, uint32_t** words_out, size_t* n_out) {
    const int saved = hash_kind();
    try {
        hash_kind() = 1;
        const Witness w = witness_stream(decode_proof(proof, len));
        hash_kind() = saved;
        std::vector<uint32_t> flat = {(uint32_t)w.vars.size(), (uint32_t)w.felts.size(), (uint32_t)w.exts.size()};
        for (auto& d : w.vars) for (int k = 0; k < 8; k++) flat.push_back(d[k].v);
        for (auto& f : w.felts) flat.push_back(f.v);
        for (auto& e : w.exts) for (int k = 0; k < 4; k++) flat.push_back(e.c[k].v);
        *words_out = (uint32_t*)malloc(flat.size() * 4 + 4);
        memcpy(*words_out, flat.data(), flat.size() * 4);
        *n_out = flat.size();
        return 0;
    } catch (const std::exception& ex) {
        hash_kind() = saved;
        g_orc_error = ex.what();
        return -1;
    }
}

}
