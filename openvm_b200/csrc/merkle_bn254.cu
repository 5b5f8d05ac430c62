// MerkleTreeMmcs of the root configuration: leaves and nodes are one BN254 scalar (eight canonical 32-bit limbs in the
// same eight digest planes the BabyBear trees use), the leaf hash is MultiField32PaddingFreeSponge<BabyBear, Bn254Fr,
// Poseidon2Bn254<3>, 3, 16, 1>, the node hash TruncatedPermutation<_, 2, 1, 3>.
// Reference twins: extensions/native/recursion/src/outer_poseidon2.rs:26-63 (p2_hash, p2_compress),
// extensions/native/recursion/src/config/outer.rs (OuterConfig types), bn254.cuh for the permutation.
//
// One thread per tree node. A permutation is 240 eight-limb Montgomery products (about 35 k integer multiply-adds), 20 x a
// BabyBear one for twice the absorbed words; the root verifier's traces are small and this path exists for completeness of
// the proof system, not for the headline commitment. The limb product and the permutation are each a single non-inlined
// device function so that a kernel holds one copy of either (instruction cache).
#include "common.cuh"
#include "bn254.cuh"

namespace b2 {

constexpr int BN_THREADS = 128;

__device__ __noinline__ void bn_permute_call(Fr* st) {
    Fr s[3] = {st[0], st[1], st[2]};
    poseidon2_bn254_permute(s);
    st[0] = s[0];
    st[1] = s[1];
    st[2] = s[2];
}

// overwrite-mode absorption of one row: blocks of 16 columns, 8 canonical words packed per element
template <class Load>
B2_D void bn_absorb_row(Fr* s, uint32_t ncols, Load load) {
#pragma unroll 1
    for (uint32_t c = 0; c < ncols; c += 16) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const uint32_t c0 = c + 8 * h;
            if (c0 < ncols) {
                uint32_t w[8];
#pragma unroll
                for (int k = 0; k < 8; k++) w[k] = (c0 + k < ncols) ? Fp::raw(load(c0 + k)).to_canonical() : 0u;
                s[h] = bn_reduce_32(w);
            }
        }
        bn_permute_call(s);
    }
}

//   mode 0: out[i] = rowhash(i)
//   mode 1: out[i] = compress(compress(prev[2i], prev[2i+1]), rowhash(i))
//   mode 2: out[i] = compress(prev[2i], prev[2i+1])
B2_D void bn_hash_node(const uint32_t* const* __restrict__ cols, uint32_t ncols, uint64_t n, const uint32_t* prev, uint32_t* out,
                       int mode, uint64_t i) {
    Fr s[3] = {fr_zero(), fr_zero(), fr_zero()};
    if (mode != 2) bn_absorb_row(s, ncols, [&](uint32_t c) { return __ldg(cols[c] + i); });
    if (mode != 0) {
        const Fr rowhash = s[0];
        uint32_t l[8], r[8];
        const uint64_t np = 2 * n;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            const uint2 lr = __ldcg(reinterpret_cast<const uint2*>(prev + (uint64_t)w * np + 2 * i));
            l[w] = lr.x;
            r[w] = lr.y;
        }
        s[0] = fr_from_words(l);
        s[1] = fr_from_words(r);
        s[2] = fr_zero();
        bn_permute_call(s);
        if (mode == 1) {
            s[1] = rowhash;
            s[2] = fr_zero();
            bn_permute_call(s);
        }
    }
    uint32_t d[8];
    fr_to_words(s[0], d);
#pragma unroll
    for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = d[w];
}

__global__ void __launch_bounds__(BN_THREADS)
    bn_tree_level_kernel(const uint32_t* const* __restrict__ cols, uint32_t ncols, uint64_t n, const uint32_t* __restrict__ prev,
                         uint32_t* __restrict__ out, int mode) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bn_hash_node(cols, ncols, n, prev, out, mode, i);
}

// streaming sponge over column groups of one matrix (the commit path hashes a group as soon as its LDE exists): the state
// of a row between groups is its three elements in 24 planes; every group but the last is a multiple of 16 columns wide
__global__ void __launch_bounds__(BN_THREADS)
    bn_leaf_absorb_kernel(const uint32_t* __restrict__ base, uint32_t ncols, uint64_t n, uint32_t* __restrict__ state,
                          uint32_t* __restrict__ out, int first, int last) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr s[3];
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int w = 0; w < 8; w++) s[k].l[w] = first ? 0u : state[(uint64_t)(8 * k + w) * n + i];
    bn_absorb_row(s, ncols, [&](uint32_t c) { return __ldg(base + (uint64_t)c * n + i); });
    if (last) {
        uint32_t d[8];
        fr_to_words(s[0], d);
#pragma unroll
        for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = d[w];
    } else {
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
            for (int w = 0; w < 8; w++) state[(uint64_t)(8 * k + w) * n + i] = s[k].l[w];
    }
}

// states: n x 24 canonical limbs (three elements each), permuted in place (b2s_poseidon2_bn254_permute)
__global__ void __launch_bounds__(BN_THREADS) bn_permute_kernel(uint32_t* st, uint64_t n) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr s[3];
    for (int k = 0; k < 3; k++) {
        uint32_t w[8];
        for (int j = 0; j < 8; j++) w[j] = st[24 * i + 8 * k + j];
        s[k] = fr_from_words(w);
    }
    bn_permute_call(s);
    for (int k = 0; k < 3; k++) {
        uint32_t w[8];
        fr_to_words(s[k], w);
        for (int j = 0; j < 8; j++) st[24 * i + 8 * k + j] = w[j];
    }
}

// proof of work under the multi-field transcript (challenger/multi_field32.rs:99-107): the witness joins the pending
// observations, one duplexing follows (at once when it completes six words, else at the sample), and the sample is the
// last output word -- limb 2 of the third scalar, mod p.
struct BnPowParams {
    uint32_t st[24];
    uint32_t pending[6];
    uint32_t n_pending, bits, base;
};
__global__ void __launch_bounds__(BN_THREADS) bn_pow_grind_kernel(BnPowParams p, uint32_t count, uint32_t* result) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const uint32_t wit = p.base + t;
    if (wit >= P) return;
    Fr s[3];
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int j = 0; j < 8; j++) s[k].l[j] = p.st[8 * k + j];
    const uint32_t n = p.n_pending + 1;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        if (3u * h < n) {
            uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const uint32_t idx = 3 * h + k;
                if (idx < n) w[k] = idx == p.n_pending ? wit : p.pending[idx];
            }
            s[h] = bn_reduce_32(w);
        }
    }
    bn_permute_call(s);
    uint32_t f[3];
    bn_split_32(s[2], f);
    if ((f[2] & ((1u << p.bits) - 1)) == 0) atomicMin(result, wit);
}

static inline unsigned bn_grid(uint64_t n) { return (unsigned)((n + BN_THREADS - 1) / BN_THREADS); }

void bn_launch_level(b2s_engine* e, const uint32_t* const* cols, uint32_t ncols, uint64_t n, const uint32_t* prev, uint32_t* out,
                     int mode) {
    bn_tree_level_kernel<<<bn_grid(n), BN_THREADS, 0, e->stream>>>(cols, ncols, n, prev, out, mode);
}
void bn_leaf_absorb(b2s_engine* e, const uint32_t* base, uint32_t ncols, uint64_t n, uint32_t* state, uint32_t* out, bool first,
                    bool last) {
    B2_REQUIRE(last || ncols % 16 == 0, B2S_ERR_SHAPE, "bn254 leaf sponge: a column group must be a multiple of 16 wide");
    bn_leaf_absorb_kernel<<<bn_grid(n), BN_THREADS, 0, e->stream>>>(base, ncols, n, state, out, first ? 1 : 0, last ? 1 : 0);
}
void bn_pow_grind(b2s_engine* e, const uint32_t state24[24], const uint32_t pending[6], uint32_t n_pending, uint32_t bits,
                  uint32_t base, uint32_t count, uint32_t* d_result) {
    B2_REQUIRE(n_pending < 6, B2S_ERR_INTERNAL, "pow: transcript buffer is full");
    BnPowParams p{};
    for (int i = 0; i < 24; i++) p.st[i] = state24[i];
    for (int i = 0; i < 6; i++) p.pending[i] = pending[i];
    p.n_pending = n_pending;
    p.bits = bits;
    p.base = base;
    bn_pow_grind_kernel<<<bn_grid(count), BN_THREADS, 0, e->stream>>>(p, count, d_result);
}
void bn_permute_states(b2s_engine* e, uint32_t* d_states, uint64_t n) {
    if (!n) return;
    bn_permute_kernel<<<bn_grid(n), BN_THREADS, 0, e->stream>>>(d_states, n);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
}

}  // namespace b2
