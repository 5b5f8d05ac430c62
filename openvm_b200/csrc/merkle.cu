// Poseidon2 Merkle commitment over column-major device matrices of mixed power-of-two heights.
//
// Replaces (reference, external): p3-merkle-tree `MerkleTreeMmcs<.., 8>::commit` as instantiated at
// extensions/native/recursion/src/hints.rs:39-46. In-tree specification of the procedure:
//   extensions/native/circuit/src/poseidon2/README.md:9-17,33-41 and the verifier's
//   extensions/native/recursion/src/fri/mod.rs:185-299 (verify_batch).
//
// Kernels (all integer-pipe bound; one Poseidon2 permutation chain per thread, state in registers):
//   row_hash_kernel   sponge over one LDE row across every matrix of a height group (coalesced: adjacent
//                     threads read adjacent rows of the same column), optionally fused with the
//                     compress-and-inject step of the level the group joins the tree at
//   compress_kernel   2-to-1 compression of one digest layer
// Digest layers are stored as 8 word-planes (word w of node i at [w*n + i]) so both kernels read and
// write fully coalesced.
#include "common.cuh"
#include "poseidon2.cuh"

namespace b2 {

constexpr int HASH_THREADS = 128;

// absorb the row `row` of `ncols` columns (pointer per column) in overwrite mode, rate 8
B2_D void sponge_row(const uint32_t* const* __restrict__ cols, uint32_t ncols, uint64_t row, Fp (&s)[16]) {
#pragma unroll
    for (int i = 0; i < 16; i++) s[i] = Fp::zero();
    uint32_t c = 0;
    for (; c + 8 <= ncols; c += 8) {
        uint32_t v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = __ldg(cols[c + k] + row);
#pragma unroll
        for (int k = 0; k < 8; k++) s[k] = Fp::raw(v[k]);
        poseidon2_permute(s);
    }
    if (c < ncols) {
        uint32_t rem = ncols - c;
#pragma unroll
        for (int k = 0; k < 8; k++)
            if ((uint32_t)k < rem) s[k] = Fp::raw(__ldg(cols[c + k] + row));
        poseidon2_permute(s);
    }
}

// mode 0: out[node] = rowhash(node)
// mode 1: out[node] = compress(compress(prev[2node], prev[2node+1]), rowhash(node))
__global__ void __launch_bounds__(HASH_THREADS) row_hash_kernel(const uint32_t* const* __restrict__ cols,
                                                                uint32_t ncols, uint64_t n,
                                                                const uint32_t* __restrict__ prev,
                                                                uint32_t* __restrict__ out, int mode) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fp s[16];
    sponge_row(cols, ncols, i, s);
    if (mode == 1) {
        Fp t[16];
        const uint64_t np = 2 * n;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            uint2 lr = *reinterpret_cast<const uint2*>(prev + (uint64_t)w * np + 2 * i);
            t[w] = Fp::raw(lr.x);
            t[8 + w] = Fp::raw(lr.y);
        }
        poseidon2_permute(t);
#pragma unroll
        for (int w = 0; w < 8; w++) { t[8 + w] = s[w]; }
        poseidon2_permute(t);
#pragma unroll
        for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = t[w].v;
    } else {
#pragma unroll
        for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = s[w].v;
    }
}

__global__ void __launch_bounds__(HASH_THREADS) compress_kernel(const uint32_t* __restrict__ prev,
                                                                uint32_t* __restrict__ out, uint64_t n) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fp t[16];
    const uint64_t np = 2 * n;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint2 lr = *reinterpret_cast<const uint2*>(prev + (uint64_t)w * np + 2 * i);
        t[w] = Fp::raw(lr.x);
        t[8 + w] = Fp::raw(lr.y);
    }
    poseidon2_permute(t);
#pragma unroll
    for (int w = 0; w < 8; w++) out[(uint64_t)w * n + i] = t[w].v;
}

static inline unsigned grid_for(uint64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

// Build the tree of d->mats (any heights). Fills d->layers.
void mmcs_build(b2s_engine* e, b2s_pcs_data* d) {
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
        std::vector<const uint32_t*> flat;
        flat.reserve(total_cols);
        for (auto& g : groups) {
            group_dev[g.first] = d_ptrs + flat.size();
            flat.insert(flat.end(), g.second.begin(), g.second.end());
        }
        B2_CUDA(cudaMemcpyAsync(d_ptrs, flat.data(), total_cols * sizeof(void*), cudaMemcpyHostToDevice, e->stream));
        B2_CUDA(cudaStreamSynchronize(e->stream));  // flat is a stack-lifetime buffer
    }
    // leaves
    d->layers.clear();
    d->layer_nodes.clear();
    {
        StageScope sc(e, ST_LEAF);
        uint32_t* leaf = (uint32_t*)e->dmalloc(hmax * 32);
        uint32_t nc = (uint32_t)groups[hmax].size();
        row_hash_kernel<<<grid_for(hmax, HASH_THREADS), HASH_THREADS, 0, e->stream>>>(group_dev[hmax], nc, hmax,
                                                                                        nullptr, leaf, 0);
        B2_CHECK_LAUNCH();
        e->launches[1]++;
        e->permutations += hmax * ((nc + 7) / 8);
        d->layers.push_back(leaf);
        d->layer_nodes.push_back(hmax);
    }
    uint64_t h = hmax;
    while (h > 1) {
        uint64_t nh = h / 2;
        uint32_t* next = (uint32_t*)e->dmalloc(nh * 32);
        auto it = groups.find(nh);
        if (it != groups.end()) {
            StageScope sc(e, ST_LEAF);
            uint32_t nc = (uint32_t)it->second.size();
            row_hash_kernel<<<grid_for(nh, HASH_THREADS), HASH_THREADS, 0, e->stream>>>(group_dev[nh], nc, nh,
                                                                                          d->layers.back(), next, 1);
            B2_CHECK_LAUNCH();
            e->launches[1]++;
            e->permutations += nh * ((nc + 7) / 8 + 2);
        } else {
            StageScope sc(e, ST_COMPRESS);
            compress_kernel<<<grid_for(nh, HASH_THREADS), HASH_THREADS, 0, e->stream>>>(d->layers.back(), next, nh);
            B2_CHECK_LAUNCH();
            e->launches[2]++;
            e->permutations += nh;
        }
        d->layers.push_back(next);
        d->layer_nodes.push_back(nh);
        h = nh;
    }
    e->dfree(d_ptrs, total_cols * sizeof(void*));
}

}  // namespace b2
