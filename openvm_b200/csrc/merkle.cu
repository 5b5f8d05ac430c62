// Poseidon2 Merkle commitment over column-major device matrices of mixed power-of-two heights.
//
// Replaces (reference, external): p3-merkle-tree `MerkleTreeMmcs<.., 8>::commit` as instantiated at
// extensions/native/recursion/src/hints.rs:39-46. In-tree specification of the procedure:
//   extensions/native/circuit/src/poseidon2/README.md:9-17,33-41 and the verifier's
//   extensions/native/recursion/src/fri/mod.rs:185-299 (verify_batch).
//
// Kernels (all integer-pipe bound; one Poseidon2 permutation chain per thread, state in registers):
//   tree_level_kernel one node per thread: sponge over one LDE row across every matrix of a height group
//                     (coalesced: adjacent threads read adjacent rows of the same column), optionally fused with
//                     the compress-and-inject step of the level the group joins the tree at; or a plain 2-to-1
//                     compression of the previous digest layer
//   rowhash_groups_kernel row digests of EVERY short height group of a commitment in one launch on a side stream, four lanes
//                     per permutation (a sponge over thousands of columns is a latency chain whatever the row count)
//   tree_subtree_kernel / tree_level_inject_kernel  the tree walk over precomputed row digests: compressions only; up to nine
//                     levels per launch below 2^14 nodes
// Digest layers are stored as 8 word-planes (word w of node i at [w*n + i]) so both kernels read and
// write fully coalesced.
#include "common.cuh"
#include "poseidon2_f64.cuh"

namespace b2 {

constexpr int HASH_THREADS = 128;
constexpr int HASH_MIN_BLOCKS = 8;  // 1024 threads/SM -> <= 64 registers/thread

// One thread produces one node of a tree level. All Poseidon2 work of a node -- the sponge over its row
// (ceil(ncols/8) permutations, overwrite mode, rate 8, no padding) and, when the row group is injected below the
// leaves, the two compressions  compress(compress(prev[2i], prev[2i+1]), rowhash(i))  -- runs through ONE
// permutation call site inside a rolled loop, so the kernel holds a single ~9 KB copy of the permutation.
// The next chunk's eight words are fetched before the current permutation starts (latency hidden behind
// ~6 k integer instructions).
//   mode 0: out[i] = rowhash(i)
//   mode 1: out[i] = compress(compress(prev[2i], prev[2i+1]), rowhash(i))
//   mode 2: out[i] = compress(prev[2i], prev[2i+1])                       (ncols == 0)
B2_D void hash_node(const uint32_t* const* __restrict__ cols, uint32_t ncols, uint64_t n,
                    const uint32_t* prev, uint32_t* out, int mode, uint64_t i) {
    Fp s[16];
#pragma unroll
    for (int k = 0; k < 16; k++) s[k] = Fp::zero();
    const uint32_t chunks = (ncols + 7) >> 3;
    const uint32_t total = chunks + (mode == 0 ? 0u : (mode == 1 ? 2u : 1u));
    uint32_t nxt[8];
    Fp keep[8];
    auto fetch = [&](uint32_t chunk) {
        const uint32_t c = chunk << 3;
#pragma unroll
        for (int k = 0; k < 8; k++) nxt[k] = (c + k < ncols) ? __ldg(cols[c + k] + i) : 0u;
    };
    if (chunks) fetch(0);
#pragma unroll 1
    for (uint32_t it = 0; it < total; it++) {
        if (it < chunks) {
            const uint32_t c = it << 3;
            if (c + 8 <= ncols) {
#pragma unroll
                for (int k = 0; k < 8; k++) s[k] = Fp::raw(nxt[k]);
            } else {
#pragma unroll
                for (int k = 0; k < 8; k++)
                    if (c + k < ncols) s[k] = Fp::raw(nxt[k]);
            }
            if (it + 1 < chunks) fetch(it + 1);
        } else if (it == chunks && mode != 0) {
            const uint64_t np = 2 * n;
#pragma unroll
            for (int w = 0; w < 8; w++) {
                keep[w] = s[w];
                // L2-coherent load: in the tail kernel `prev` was written by this block one barrier earlier
                uint2 lr = __ldcg(reinterpret_cast<const uint2*>(prev + (uint64_t)w * np + 2 * i));
                s[w] = Fp::raw(lr.x);
                s[8 + w] = Fp::raw(lr.y);
            }
        } else {
#pragma unroll
            for (int w = 0; w < 8; w++) s[8 + w] = keep[w];
        }
        poseidon2_permute(s);
    }
#pragma unroll
    for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = s[w].v;
}

__global__ void __launch_bounds__(HASH_THREADS, HASH_MIN_BLOCKS)
    tree_level_kernel(const uint32_t* const* __restrict__ cols, uint32_t ncols, uint64_t n,
                      const uint32_t* __restrict__ prev, uint32_t* __restrict__ out, int mode) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    hash_node(cols, ncols, n, prev, out, mode, i);
}

// compress(prev[2i], prev[2i+1]), then compress(that, inj[i]): a level above SUBTREE_MAX_FIRST nodes whose row digests were
// computed by rowhash_groups_kernel (one node per thread)
__global__ void __launch_bounds__(HASH_THREADS, HASH_MIN_BLOCKS)
    tree_level_inject_kernel(const uint32_t* __restrict__ prev, const uint32_t* __restrict__ inj, uint32_t* __restrict__ out, uint64_t n) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fp s[16];
    const uint64_t np = 2 * n;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const uint2 lr = __ldcg(reinterpret_cast<const uint2*>(prev + (uint64_t)w * np + 2 * i));
        s[w] = Fp::raw(lr.x);
        s[8 + w] = Fp::raw(lr.y);
    }
    poseidon2_permute(s);
#pragma unroll
    for (int w = 0; w < 8; w++) s[8 + w] = Fp::raw(__ldg(inj + (uint64_t)w * n + i));
    poseidon2_permute(s);
#pragma unroll
    for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = s[w].v;
}

// ---- four lanes per permutation --------------------------------------------------------------------------------------
// Small levels (the top of every tree, every level of the FRI commit-phase trees) are bound by the LATENCY of one
// permutation chain per thread (~8 us: 5.7 k dependent-ish instructions), not by throughput. Here four adjacent lanes
// share one state, lane `sub` holding words 4*sub .. 4*sub+3: the S-box layer and the 4x4 blocks of the external layer
// stay lane-local, the column sums of the external layer and the row sum of the internal layer are two xor-shuffles
// each, the internal diagonal is one Montgomery product per word with per-lane constants. ~2 us per permutation.
static __device__ uint32_t P2G_EXT[8][16] = B2_P2_EXT_ALL;
static __device__ uint32_t P2G_DIAG[16] = B2_P2_DIAG;
constexpr unsigned FULL_MASK = 0xFFFFFFFFu;

B2_D Fp quad_sum(Fp v) {
    v = v + Fp::raw(__shfl_xor_sync(FULL_MASK, v.v, 1));
    return v + Fp::raw(__shfl_xor_sync(FULL_MASK, v.v, 2));
}
B2_D void poseidon2_permute4(Fp (&s)[4], const Fp (&diag)[4], uint32_t sub) {
#pragma unroll 1
    for (int r = 0; r < 9; r++) {
        uint4 rc = make_uint4(0, 0, 0, 0);
        if (r < 8) rc = __ldg(reinterpret_cast<const uint4*>(&P2G_EXT[r][4 * sub]));
        mat4(s[0], s[1], s[2], s[3]);
#pragma unroll
        for (int k = 0; k < 4; k++) s[k] += quad_sum(s[k]);
        if (r == 4) {
#pragma unroll 1
            for (int q = 0; q < 13; q++) {
                const Fp x0 = sbox7_rc(s[0], P2C_PARTIAL[q]);
                if (sub == 0) s[0] = x0;
                const Fp sum = quad_sum((s[0] + s[1]) + (s[2] + s[3]));
#pragma unroll
                for (int k = 0; k < 4; k++) s[k] = s[k] * diag[k] + sum;
            }
        }
        if (r < 8) {
            s[0] = sbox7_rc(s[0], rc.x);
            s[1] = sbox7_rc(s[1], rc.y);
            s[2] = sbox7_rc(s[2], rc.z);
            s[3] = sbox7_rc(s[3], rc.w);
        }
    }
}
// hash_node with four lanes per node; every lane of the warp must call it (idle quads pass store = false and a valid i)
B2_D void hash_node4(const uint32_t* const* __restrict__ cols, uint32_t ncols, uint64_t n, const uint32_t* prev,
                     uint32_t* out, int mode, uint64_t i, uint32_t sub, bool store) {
    Fp s[4], keep[4], diag[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        s[k] = Fp::zero();
        keep[k] = Fp::zero();
        diag[k] = Fp::raw(P2G_DIAG[4 * sub + k]);
    }
    const uint32_t chunks = (ncols + 7) >> 3;
    const uint32_t total = chunks + (mode == 0 ? 0u : (mode == 1 ? 2u : 1u));
    const unsigned lane = threadIdx.x & 31u;
    // the next chunk's words are fetched before the current permutation starts: two dependent loads (pointer table, column)
    // otherwise sit between every two permutations of a long row (a 3164-column row is 396 of them)
    uint32_t nxt[4] = {0, 0, 0, 0};
    auto fetch = [&](uint32_t it) {
        const uint32_t c = (it << 3) + 4 * sub;
        if (sub < 2) {
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (c + k < ncols) nxt[k] = __ldg(cols[c + k] + i);
        }
    };
    if (chunks) fetch(0);
#pragma unroll 1
    for (uint32_t it = 0; it < total; it++) {
        if (it < chunks) {
            const uint32_t c = (it << 3) + 4 * sub;
            if (sub < 2) {
#pragma unroll
                for (int k = 0; k < 4; k++)
                    if (c + k < ncols) s[k] = Fp::raw(nxt[k]);
            }
            if (it + 1 < chunks) fetch(it + 1);
        } else if (it == chunks && mode != 0) {
            const uint64_t np = 2 * n, node = 2 * i + (sub >> 1);
            const uint32_t w0 = 4 * (sub & 1);
#pragma unroll
            for (int k = 0; k < 4; k++) {
                keep[k] = s[k];
                s[k] = Fp::raw(__ldcg(prev + (uint64_t)(w0 + k) * np + node));
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t v = __shfl_sync(FULL_MASK, keep[k].v, sub >= 2 ? lane - 2 : lane);
                if (sub >= 2) s[k] = Fp::raw(v);
            }
        }
        poseidon2_permute4(s, diag, sub);
    }
    if (store && sub < 2) {
#pragma unroll
        for (int k = 0; k < 4; k++) out[(uint64_t)(4 * sub + k) * n + i] = s[k].v;
    }
}

// compress(prev[2i], prev[2i+1]) and, when the level has row digests to inject, compress(that, inj[i]); four lanes per node
// as in hash_node4 (lane `sub` holds state words 4 sub .. 4 sub + 3). The row digests come from rowhash_groups_kernel.
B2_D void compress_node4(const uint32_t* prev, const uint32_t* inj, uint32_t* out, uint64_t n, uint64_t i, uint32_t sub, bool store) {
    Fp s[4], diag[4];
    const uint64_t np = 2 * n, node = 2 * i + (sub >> 1);
    const uint32_t w0 = 4 * (sub & 1);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        diag[k] = Fp::raw(P2G_DIAG[4 * sub + k]);
        s[k] = Fp::raw(__ldcg(prev + (uint64_t)(w0 + k) * np + node));
    }
    poseidon2_permute4(s, diag, sub);
    if (inj) {
        if (sub >= 2) {
#pragma unroll
            for (int k = 0; k < 4; k++) s[k] = Fp::raw(__ldcg(inj + (uint64_t)(w0 + k) * n + i));
        }
        poseidon2_permute4(s, diag, sub);
    }
    if (store && sub < 2) {
#pragma unroll
        for (int k = 0; k < 4; k++) out[(uint64_t)(4 * sub + k) * n + i] = s[k].v;
    }
}

// Row digests of EVERY height group of a commitment in one launch (groups of at most ROWHASH_MAX_ROWS rows): a sponge over a
// 3164-column row is 396 dependent permutations (~1 ms) whatever the row count, and a segment's tree has a dozen such groups;
// hashed level by level inside the tree walk their latencies added up (r2h launch list: tree_subtree_kernel 9 ms of a
// 44 ms proof). Here they overlap, and the tree walk only compresses.
constexpr int ROWHASH_MAX_GROUPS = 32;
constexpr uint64_t ROWHASH_MAX_ROWS = 1u << 18;
constexpr uint64_t ROWHASH_MAX_PERMS = 1u << 20;   // above SUBTREE_MAX_FIRST rows only groups this cheap join the launch
constexpr int ROWHASH_THREADS = 256;
struct RowHashParams {
    const uint32_t* const* cols[ROWHASH_MAX_GROUPS];
    uint32_t* out[ROWHASH_MAX_GROUPS];
    uint32_t ncols[ROWHASH_MAX_GROUPS];
    uint32_t n[ROWHASH_MAX_GROUPS];
    uint32_t first_block[ROWHASH_MAX_GROUPS + 1];
    uint32_t ngroups;
};
__global__ void __launch_bounds__(ROWHASH_THREADS) rowhash_groups_kernel(RowHashParams p) {
    uint32_t g = 0;
    while (g + 1 < p.ngroups && blockIdx.x >= p.first_block[g + 1]) g++;
    const uint64_t n = p.n[g];
    const uint64_t t = (uint64_t)(blockIdx.x - p.first_block[g]) * blockDim.x + threadIdx.x;
    const uint64_t i = t >> 2;
    // whole warps past the end leave together; inside a warp idle quads keep the shuffles converged
    if ((t & ~31ull) >> 2 >= n) return;
    hash_node4(p.cols[g], p.ncols[g], n, nullptr, p.out[g], 0, i < n ? i : n - 1, (uint32_t)(t & 3), i < n);
}

// Several consecutive levels in one launch: block b owns the sub-tree over nodes [b*W, (b+1)*W) of the first level
// (W = min(n0, SUBTREE_NODES)) and walks it up to its single top node with a block barrier per level, four lanes per
// node, instead of one launch (and one ~8 us latency-bound permutation wave) per level. With one block it is the top of
// the tree; with n0 / W blocks it takes log2(W) + 1 levels off a mid-sized tree.
constexpr int SUBTREE_NODES = 256;
constexpr int SUBTREE_THREADS = 4 * SUBTREE_NODES;
constexpr int SUBTREE_MAX_LEVELS = 9;
constexpr uint64_t SUBTREE_MAX_FIRST = 1u << 14;  // one wave of blocks; above this the levels run one launch each
struct SubtreeParams {
    const uint32_t* inj[SUBTREE_MAX_LEVELS];   // row digests injected at level l (rowhash_groups_kernel), or null
    uint32_t* out[SUBTREE_MAX_LEVELS];
    const uint32_t* prev0;
    uint64_t n0;        // nodes of the first level handled here
    uint32_t w0;        // nodes of the first level per block
    uint32_t levels;
};
__global__ void __launch_bounds__(SUBTREE_THREADS, 1) tree_subtree_kernel(SubtreeParams p) {
    const uint32_t* prev = p.prev0;
    uint64_t n = p.n0;
    uint32_t w = p.w0;
    const uint32_t local = threadIdx.x >> 2, sub = threadIdx.x & 3;
#pragma unroll 1
    for (uint32_t l = 0; l < p.levels; l++) {
        // whole warps drop out together (8 nodes per warp): the shuffles stay converged
        if ((threadIdx.x >> 5) * 8 < w) {
            const bool valid = local < w;
            const uint64_t i = (uint64_t)blockIdx.x * w + (valid ? local : w - 1);
            compress_node4(prev, p.inj[l], p.out[l], n, i, sub, valid);
        }
        __syncthreads();
        prev = p.out[l];
        n >>= 1;
        w >>= 1;
    }
}

// Streaming leaf sponge for a tall matrix whose columns arrive in groups (commit_host_streamed, api.cu): absorbs
// columns [0, ncols) of `base` (column stride = n) into the per-row sponge states. `state` holds the 16 words of every
// row as word-planes between groups (read unless `first`, written unless `last`); the last group writes the digest
// planes to `out`. Every group but the last has ncols % 8 == 0, so a group boundary is always a permutation boundary.
__global__ void __launch_bounds__(HASH_THREADS, HASH_MIN_BLOCKS)
    leaf_absorb_kernel(const uint32_t* __restrict__ base, uint32_t ncols, uint64_t n, uint32_t* __restrict__ state,
                       uint32_t* __restrict__ out, int first, int last) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fp s[16];
#pragma unroll
    for (int k = 0; k < 16; k++) s[k] = first ? Fp::zero() : Fp::raw(state[(uint64_t)k * n + i]);
    const uint32_t chunks = (ncols + 7) >> 3;
    uint32_t nxt[8];
    auto fetch = [&](uint32_t chunk) {
        const uint32_t c = chunk << 3;
#pragma unroll
        for (int k = 0; k < 8; k++) nxt[k] = (c + k < ncols) ? __ldg(base + (uint64_t)(c + k) * n + i) : 0u;
    };
    if (chunks) fetch(0);
#pragma unroll 1
    for (uint32_t it = 0; it < chunks; it++) {
        const uint32_t c = it << 3;
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (c + k < ncols) s[k] = Fp::raw(nxt[k]);
        if (it + 1 < chunks) fetch(it + 1);
        poseidon2_permute(s);
    }
    if (last) {
#pragma unroll
        for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = s[w].v;
    } else {
#pragma unroll
        for (int k = 0; k < 16; k++) state[(uint64_t)k * n + i] = s[k].v;
    }
}

// ---- integer and FP64 pipes side by side ---------------------------------------------------------------------------
// Tall leaf levels (the bulk of every commit) run as a persistent kernel whose warps take tiles of 32 adjacent rows from
// an atomic counter. MIX_INT_WARPS of the 8 warps of a block hash their rows with the integer permutation, the others
// with the FP64 one (poseidon2_f64.cuh): the fmaheavy pipe (IMAD) and the FP64 pipe (DFMA) work at the same time, and a
// tile's digest words do not depend on which code path produced them (both are exact arithmetic mod p). Which tile goes
// to which path is decided by whoever is free, so the faster path simply takes more tiles.
constexpr int MIX_THREADS = 256;
constexpr int MIX_INT_WARPS = 4;
constexpr uint64_t MIX_MIN_ROWS = 1u << 19;  // >= 4 tiles per resident warp; below, a level is bound by the slower (FP64) path's latency

struct ColTable {     // columns through a pointer table (any mix of matrices of one height)
    const uint32_t* const* cols;
    B2_D uint32_t ld(uint32_t c, uint64_t i) const { return __ldg(cols[c] + i); }
};
struct ColStrided {   // columns of one matrix
    const uint32_t* base;
    uint64_t stride;
    B2_D uint32_t ld(uint32_t c, uint64_t i) const { return __ldg(base + (uint64_t)c * stride + i); }
};

// sponge over columns [0, ncols) of row i. state_in: 16 word-planes to resume from (or null: zero state);
// state_out: 16 word-planes to suspend to (or null: write the 8 digest planes to out)
// One node: sponge over columns [0, ncols) of row i, then (PREV) the compressions that join the level below:
//   node = compress(compress(prev[2i], prev[2i+1]), rowhash)   or, with ncols == 0, the inner compression alone.
// state_in: 16 word-planes to resume from (or null: zero state); state_out: 16 word-planes to suspend to (or null: write
// the 8 digest planes to out). As in hash_node, every permutation of a node runs through ONE call site inside a rolled
// loop: with three call sites per code path the kernel's ~60 KB of permutation bodies miss the instruction cache
// (measured: the leaf-aggregation shape's leaf stage 210 -> 237 ms).
template <bool PREV, class Cols>
B2_D void sponge_row_int(const Cols& src, uint32_t ncols, uint64_t n, uint64_t i, const uint32_t* state_in,
                         uint32_t* state_out, uint32_t* out, const uint32_t* prev) {
    Fp s[16];
#pragma unroll
    for (int k = 0; k < 16; k++) s[k] = state_in ? Fp::raw(state_in[(uint64_t)k * n + i]) : Fp::zero();
    const uint32_t chunks = (ncols + 7) >> 3;
    const uint32_t total = chunks + (PREV ? (ncols ? 2u : 1u) : 0u);
    uint32_t nxt[8];
    Fp keep[PREV ? 8 : 1];
    auto fetch = [&](uint32_t chunk) {
        const uint32_t c = chunk << 3;
#pragma unroll
        for (int k = 0; k < 8; k++) nxt[k] = (c + k < ncols) ? src.ld(c + k, i) : 0u;
    };
    if (chunks) fetch(0);
#pragma unroll 1
    for (uint32_t it = 0; it < total; it++) {
        if (!PREV || it < chunks) {
            const uint32_t c = it << 3;
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (c + k < ncols) s[k] = Fp::raw(nxt[k]);
            if (it + 1 < chunks) fetch(it + 1);
        } else if (it == chunks) {
            const uint64_t np = 2 * n;
#pragma unroll
            for (int w = 0; w < 8; w++) {
                keep[w] = s[w];
                const uint2 lr = __ldcg(reinterpret_cast<const uint2*>(prev + (uint64_t)w * np + 2 * i));
                s[w] = Fp::raw(lr.x);
                s[8 + w] = Fp::raw(lr.y);
            }
        } else {
#pragma unroll
            for (int w = 0; w < 8; w++) s[8 + w] = keep[w];
        }
        poseidon2_permute(s);
    }
    if (state_out) {
#pragma unroll
        for (int k = 0; k < 16; k++) state_out[(uint64_t)k * n + i] = s[k].v;
    } else {
#pragma unroll
        for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = s[w].v;
    }
}
template <bool PREV, class Cols>
B2_D void sponge_row_f64(const Cols& src, uint32_t ncols, uint64_t n, uint64_t i, const uint32_t* state_in,
                         uint32_t* state_out, uint32_t* out, const uint32_t* prev) {
    double s[16];
#pragma unroll
    for (int k = 0; k < 16; k++) s[k] = state_in ? f64_from_monty(state_in[(uint64_t)k * n + i]) : 0.0;
    const uint32_t chunks = (ncols + 7) >> 3;
    const uint32_t total = chunks + (PREV ? (ncols ? 2u : 1u) : 0u);
    uint32_t nxt[8];
    double keep[PREV ? 8 : 1];
    auto fetch = [&](uint32_t chunk) {
        const uint32_t c = chunk << 3;
#pragma unroll
        for (int k = 0; k < 8; k++) nxt[k] = (c + k < ncols) ? src.ld(c + k, i) : 0u;
    };
    if (chunks) fetch(0);
#pragma unroll 1
    for (uint32_t it = 0; it < total; it++) {
        if (!PREV || it < chunks) {
            const uint32_t c = it << 3;
#pragma unroll
            for (int k = 0; k < 8; k++)
                if (c + k < ncols) s[k] = f64_from_monty(nxt[k]);
            if (it + 1 < chunks) fetch(it + 1);
        } else if (it == chunks) {
            const uint64_t np = 2 * n;
#pragma unroll
            for (int w = 0; w < 8; w++) {
                keep[w] = s[w];
                const uint2 lr = __ldcg(reinterpret_cast<const uint2*>(prev + (uint64_t)w * np + 2 * i));
                s[w] = f64_from_monty(lr.x);
                s[8 + w] = f64_from_monty(lr.y);
            }
        } else {
#pragma unroll
            for (int w = 0; w < 8; w++) s[8 + w] = keep[w];
        }
        poseidon2_permute_f64(s);   // words stay within 17.5 p between permutations, no reduction needed
    }
    if (state_out) {
#pragma unroll
        for (int k = 0; k < 16; k++) state_out[(uint64_t)k * n + i] = f64_to_monty(s[k]);
    } else {
#pragma unroll
        for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = f64_to_monty(s[w]);
    }
}
// n is a multiple of 32 (power of two >= MIX_MIN_ROWS). ctr: one word, zeroed on the stream before the launch.
// 3 blocks of 8 warps per SM (80 registers): measured against 4 x 8 warps at 64 registers, 8 x 4 warps, and an
// all-integer plus an all-FP64 kernel side by side on two streams -- all end on the same plateau (236-239 ms for the
// 2^25 x 256 leaf level, 259 ms integer only), which is the issue rate of the two code paths together, not a pipe.
template <class Cols, bool PREV>
__global__ void __launch_bounds__(MIX_THREADS, 3) leaf_mixed_kernel(Cols src, uint32_t ncols, uint64_t n, const uint32_t* state_in,
                                                                    uint32_t* state_out, uint32_t* out, const uint32_t* prev, unsigned* ctr,
                                                                    int int_warps) {
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tiles = (unsigned)(n >> 5);
    const bool int_path = (int)warp < int_warps;
    // the next tile is claimed before the current one is hashed: the counter's round trip (~1 us) hides behind the
    // permutations even on levels with one or two of them per row
    unsigned t = 0;
    if (lane == 0) t = atomicAdd(ctr, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    while (t < tiles) {
        unsigned nxt_t = 0;
        if (lane == 0) nxt_t = atomicAdd(ctr, 1u);
        const uint64_t i = ((uint64_t)t << 5) + lane;
        if (int_path) sponge_row_int<PREV>(src, ncols, n, i, state_in, state_out, out, prev);
        else sponge_row_f64<PREV>(src, ncols, n, i, state_in, state_out, out, prev);
        t = __shfl_sync(0xffffffffu, nxt_t, 0);
    }
}

// B2S_HASH_F64: number of the 8 warps per block left on the integer path (default MIX_INT_WARPS; 8 = all integer
// through the persistent kernel; "off" = the one-row-per-thread integer kernels)
static uint64_t mixed_min_rows() {  // B2S_HASH_MIN_LOG: log2 of the smallest level hashed by the two-pipe kernel
    static uint64_t v = [] {
        const char* env = getenv("B2S_HASH_MIN_LOG");
        if (!env) return MIX_MIN_ROWS;
        const int l = atoi(env);
        return (uint64_t)1 << (l < 5 ? 5 : (l > 40 ? 40 : l));   // a tile is 32 rows
    }();
    return v;
}
static int mixed_int_warps() {
    static int v = [] {
        const char* env = getenv("B2S_HASH_F64");
        if (!env) return MIX_INT_WARPS;
        if (!strcmp(env, "off")) return -1;
        const int k = atoi(env);
        return k < 0 ? 0 : (k > 8 ? 8 : k);
    }();
    return v;
}
template <class Cols, bool PREV>
static void launch_mixed_as(b2s_engine* e, Cols src, uint32_t ncols, uint64_t n, const uint32_t* state_in, uint32_t* state_out,
                            uint32_t* out, const uint32_t* prev) {
    // resident blocks of the whole chip, computed once (thread-safe: engines on several host threads share this)
    static const unsigned resident = [] {
        int per_sm = 0, dev = 0, sms = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, leaf_mixed_kernel<Cols, PREV>, MIX_THREADS, 0) != cudaSuccess ||
            cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
            return 0u;
        return (unsigned)(sms * (per_sm < 1 ? 1 : per_sm));
    }();
    B2_REQUIRE(resident > 0, B2S_ERR_INTERNAL, "leaf hash: occupancy query failed");
    if (!e->tile_ctr) e->tile_ctr = (unsigned*)e->dmalloc(256);
    B2_CUDA(cudaMemsetAsync(e->tile_ctr, 0, 4, e->stream));
    const uint64_t want = (n >> 5) / (MIX_THREADS / 32) + 1;  // no more blocks than tiles / warps
    const unsigned grid = (unsigned)std::min<uint64_t>(resident, want);
    leaf_mixed_kernel<Cols, PREV><<<grid, MIX_THREADS, 0, e->stream>>>(src, ncols, n, state_in, state_out, out, prev, e->tile_ctr,
                                                                       mixed_int_warps());
}
template <class Cols>
static void launch_mixed(b2s_engine* e, Cols src, uint32_t ncols, uint64_t n, const uint32_t* state_in, uint32_t* state_out,
                         uint32_t* out, const uint32_t* prev = nullptr) {
    if (prev) launch_mixed_as<Cols, true>(e, src, ncols, n, state_in, state_out, out, prev);
    else launch_mixed_as<Cols, false>(e, src, ncols, n, state_in, state_out, out, prev);
}

static inline unsigned grid_for(uint64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

void leaf_absorb(b2s_engine* e, const uint32_t* base, uint32_t ncols, uint64_t n, uint32_t* state, uint32_t* out, bool first,
                 bool last) {
    StageScope sc(e, ST_LEAF);
    if (e->hash_kind) {
        bn_leaf_absorb(e, base, ncols, n, state, out, first, last);
        B2_CHECK_LAUNCH();
        e->launches[1]++;
        e->permutations += n * ((ncols + 15) / 16);
        return;
    }
    if (n >= mixed_min_rows() && mixed_int_warps() >= 0)
        launch_mixed(e, ColStrided{base, n}, ncols, n, first ? nullptr : state, last ? nullptr : state, out);
    else
        leaf_absorb_kernel<<<grid_for(n, HASH_THREADS), HASH_THREADS, 0, e->stream>>>(base, ncols, n, state, out, first ? 1 : 0,
                                                                                       last ? 1 : 0);
    B2_CHECK_LAUNCH();
    e->launches[1]++;
    e->permutations += n * ((ncols + 7) / 8);
}

// one tree level above the multi-level kernel's range: the two-pipe persistent kernel from 2^19 nodes, else one node per thread
static void launch_level(b2s_engine* e, const uint32_t* const* cols, uint32_t ncols, uint64_t n, const uint32_t* prev,
                         uint32_t* out, int mode) {
    if (e->hash_kind)
        bn_launch_level(e, cols, ncols, n, prev, out, mode);
    else if (n >= mixed_min_rows() && mixed_int_warps() >= 0)
        launch_mixed(e, ColTable{cols}, ncols, n, nullptr, nullptr, out, mode == 0 ? nullptr : prev);
    else
        tree_level_kernel<<<grid_for(n, HASH_THREADS), HASH_THREADS, 0, e->stream>>>(cols, ncols, n, prev, out, mode);
}

// Build the tree of d->mats (any heights). Fills d->layers. `leaf_layer` (optional): digest planes of the tallest height
// group, already computed by the streaming sponge; the tree takes ownership of it.
void mmcs_build(b2s_engine* e, b2s_pcs_data* d, uint32_t* leaf_layer) {
    B2_REQUIRE(!d->mats.empty(), B2S_ERR_BAD_ARG, "mmcs: no matrices");
    uint64_t hmax = 0;
    for (auto* m : d->mats) {
        B2_REQUIRE(is_pow2(m->height), B2S_ERR_SHAPE, "mmcs: matrix height must be a power of two");
        B2_REQUIRE(m->width > 0, B2S_ERR_SHAPE, "mmcs: zero-width matrix");
        if (m->height > hmax) hmax = m->height;
    }
    d->log_max_height = log2_exact(hmax);
    // column pointer lists per height, matrices of equal height in caller order (stable sort by height)
    std::map<uint64_t, std::vector<const uint32_t*>> groups;
    for (auto* m : d->mats)
        for (uint64_t c = 0; c < m->width; c++) groups[m->height].push_back(m->d + c * m->height);
    size_t total_cols = 0;
    for (auto& g : groups) total_cols += g.second.size();
    const uint32_t** d_ptrs = (const uint32_t**)e->dmalloc(total_cols * sizeof(void*));
    std::map<uint64_t, const uint32_t**> group_dev;
    {
        // the host table lives in the PcsData: no stream round trip is needed before it may go out of scope
        std::vector<const uint32_t*>& flat = d->col_ptrs;
        flat.clear();
        flat.reserve(total_cols);
        for (auto& g : groups) {
            group_dev[g.first] = d_ptrs + flat.size();
            flat.insert(flat.end(), g.second.begin(), g.second.end());
        }
        B2_CUDA(cudaMemcpyAsync(d_ptrs, flat.data(), total_cols * sizeof(void*), cudaMemcpyHostToDevice, e->stream));
    }
    d->layers.clear();
    d->layer_nodes.clear();
    // ---- row digests of every group below the two-pipe kernel's threshold, one launch (Poseidon2-BabyBear only)
    std::map<uint64_t, uint32_t*> inj;           // height -> digest planes (the tallest group's go straight into layer 0)
    std::vector<std::pair<uint32_t*, size_t>> inj_bufs;
    uint32_t* small_leaf = nullptr;
    cudaEvent_t rowhash_done = nullptr;
    if (!e->hash_kind) {
        RowHashParams rp{};
        uint32_t blocks = 0;
        for (auto it = groups.rbegin(); it != groups.rend(); ++it) {   // tallest first
            const uint64_t h = it->first;
            // every group the multi-level kernel meets (<= SUBTREE_MAX_FIRST rows) needs its digests; taller ones up to the
            // two-pipe kernel's threshold join the launch so that their sponges overlap the others
            if (h == hmax && leaf_layer) continue;
            // ... unless they are real work: the four-lane sponge spends ~1.5x the multiplies of the one-row-per-thread kernel,
            // so a tall AND wide group (KeccakVmAir shape: 2^18 x 3164, main commit 33.6 -> 48.8 ms) stays on the throughput path
            const uint64_t perms = h * ((it->second.size() + 7) / 8);
            if (h > SUBTREE_MAX_FIRST && (h > ROWHASH_MAX_ROWS || h >= mixed_min_rows() || perms > ROWHASH_MAX_PERMS)) continue;
            B2_REQUIRE(rp.ngroups < ROWHASH_MAX_GROUPS, B2S_ERR_SHAPE, "mmcs: too many distinct matrix heights");
            uint32_t* out = (uint32_t*)e->dmalloc(h * 32);
            if (h == hmax) small_leaf = out;
            else { inj[h] = out; inj_bufs.push_back({out, h * 32}); }
            const uint32_t g = rp.ngroups++;
            rp.cols[g] = group_dev[h];
            rp.ncols[g] = (uint32_t)it->second.size();
            rp.n[g] = (uint32_t)h;
            rp.out[g] = out;
            rp.first_block[g] = blocks;
            blocks += (uint32_t)((4 * h + ROWHASH_THREADS - 1) / ROWHASH_THREADS);
            e->permutations += h * ((rp.ncols[g] + 7) / 8);
        }
        rp.first_block[rp.ngroups] = blocks;
        if (rp.ngroups) {
            // On a side stream: the launch is as long as its longest row (milliseconds for thousands of columns) and uses a
            // few SMs; the tall levels (two-pipe leaf kernel, big compress levels) run next to it on the engine's stream.
            if (!e->hash_stream) B2_CUDA(cudaStreamCreateWithFlags(&e->hash_stream, cudaStreamNonBlocking));
            cudaEvent_t fork = e->get_event();
            B2_CUDA(cudaEventRecord(fork, e->stream));
            B2_CUDA(cudaStreamWaitEvent(e->hash_stream, fork, 0));
            e->event_pool.push_back(fork);
            rowhash_groups_kernel<<<blocks, ROWHASH_THREADS, 0, e->hash_stream>>>(rp);
            B2_CHECK_LAUNCH();
            e->launches[1]++;
            rowhash_done = e->get_event();
            B2_CUDA(cudaEventRecord(rowhash_done, e->hash_stream));
        }
    }
    auto need_rowhash = [&]() {   // first reader of the row digests on the engine's stream
        if (!rowhash_done) return;
        B2_CUDA(cudaStreamWaitEvent(e->stream, rowhash_done, 0));
        e->event_pool.push_back(rowhash_done);
        rowhash_done = nullptr;
    };
    // leaves
    if (leaf_layer || small_leaf) {
        if (small_leaf) need_rowhash();
        d->layers.push_back(leaf_layer ? leaf_layer : small_leaf);
        d->layer_nodes.push_back(hmax);
    } else {
        StageScope sc(e, ST_LEAF);
        uint32_t* leaf = (uint32_t*)e->dmalloc(hmax * 32);
        uint32_t nc = (uint32_t)groups[hmax].size();
        launch_level(e, group_dev[hmax], nc, hmax, nullptr, leaf, 0);
        B2_CHECK_LAUNCH();
        e->launches[1]++;
        e->permutations += hmax * ((nc + 7) / 8);
        d->layers.push_back(leaf);
        d->layer_nodes.push_back(hmax);
    }
    uint64_t h = hmax;
    while (h > 1) {
        uint64_t nh = h / 2;
        if (!e->hash_kind && nh <= SUBTREE_MAX_FIRST) {
            // log2(W) + 1 levels in one launch (all remaining ones when nh <= SUBTREE_NODES): compressions only, the row
            // digests of the groups that join on the way are already there
            StageScope sc(e, ST_COMPRESS);
            SubtreeParams tp{};
            tp.prev0 = d->layers.back();
            tp.n0 = nh;
            tp.w0 = (uint32_t)std::min<uint64_t>(nh, SUBTREE_NODES);
            uint32_t l = 0;
            uint64_t m = nh;
            for (uint32_t w = tp.w0; w >= 1; w >>= 1, m >>= 1, l++) {
                auto it = inj.find(m);
                tp.inj[l] = it != inj.end() ? it->second : nullptr;
                tp.out[l] = (uint32_t*)e->dmalloc(m * 32);
                d->layers.push_back(tp.out[l]);
                d->layer_nodes.push_back(m);
                e->permutations += m * (tp.inj[l] ? 2 : 1);
            }
            tp.levels = l;
            need_rowhash();
            tree_subtree_kernel<<<(unsigned)(nh / tp.w0), SUBTREE_THREADS, 0, e->stream>>>(tp);
            B2_CHECK_LAUNCH();
            e->launches[2]++;
            h = nh / tp.w0;   // nodes of the last level written
            continue;
        }
        uint32_t* next = (uint32_t*)e->dmalloc(nh * 32);
        auto it = groups.find(nh);
        auto pre = inj.find(nh);
        if (pre != inj.end()) {
            need_rowhash();
            StageScope sc(e, ST_COMPRESS);
            tree_level_inject_kernel<<<grid_for(nh, HASH_THREADS), HASH_THREADS, 0, e->stream>>>(d->layers.back(), pre->second, next, nh);
            B2_CHECK_LAUNCH();
            e->launches[2]++;
            e->permutations += 2 * nh;
        } else if (it != groups.end()) {
            StageScope sc(e, ST_LEAF);
            uint32_t nc = (uint32_t)it->second.size();
            launch_level(e, group_dev[nh], nc, nh, d->layers.back(), next, 1);
            B2_CHECK_LAUNCH();
            e->launches[1]++;
            e->permutations += nh * ((nc + 7) / 8 + 2);
        } else {
            StageScope sc(e, ST_COMPRESS);
            launch_level(e, nullptr, 0, nh, d->layers.back(), next, 2);
            B2_CHECK_LAUNCH();
            e->launches[2]++;
            e->permutations += nh;
        }
        d->layers.push_back(next);
        d->layer_nodes.push_back(nh);
        h = nh;
    }
    need_rowhash();   // (nothing read them: a one-row tree) the buffers are freed on the engine's stream
    for (auto& b : inj_bufs) e->dfree(b.first, b.second);
    e->dfree(d_ptrs, total_cols * sizeof(void*));
}

}  // namespace b2
