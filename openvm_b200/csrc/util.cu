// Data-movement and self-test kernels: representation conversion, row-major <-> column-major transposes
// (DeviceDataTransporter; crates/vm/src/arch/vm.rs:731-734,1307-1312), synthetic fill, Merkle opening
// gather (Mmcs::open_batch; verifier side extensions/native/recursion/src/fri/mod.rs:185-299), and the
// integer-pipe micro-benchmarks that give the modmul/s and permutations/s "peaks" quoted in DESIGN.md.
#include "common.cuh"
#include "poseidon2.cuh"

namespace b2 {

__global__ void permute_states_kernel(uint32_t* st, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fp s[16];
#pragma unroll
    for (int k = 0; k < 16; k++) s[k] = Fp::from_canonical(st[16 * i + k]);
    poseidon2_permute(s);
#pragma unroll
    for (int k = 0; k < 16; k++) st[16 * i + k] = s[k].to_canonical();
}
void poseidon2_permute_states(b2s_engine* e, uint32_t* d_states, size_t n) {
    if (!n) return;
    permute_states_kernel<<<(unsigned)((n + 127) / 128), 128, 0, e->stream>>>(d_states, n);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
    e->permutations += n;
}

__global__ void convert_kernel(uint32_t* d, size_t n, int to_monty) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) d[i] = to_monty ? Fp::from_canonical(d[i]).v : Fp::raw(d[i]).to_canonical();
}
void convert_repr(b2s_engine* e, uint32_t* d, size_t n, bool to_monty) {
    if (!n) return;
    unsigned blocks = (unsigned)std::min<size_t>((n + 255) / 256, (size_t)e->num_sms * 16);
    convert_kernel<<<blocks, 256, 0, e->stream>>>(d, n, to_monty ? 1 : 0);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
}

// 32x32 tiled transpose through shared memory; conv: 0 none, 1 canonical->monty, 2 monty->canonical.
// in is `rows_in x cols_in` with row stride cols_in; out is `cols_in x rows_in`.
__global__ void transpose_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t rows_in,
                                 uint64_t cols_in, int conv) {
    __shared__ uint32_t tile[32][33];
    uint64_t bx = (uint64_t)blockIdx.x * 32, by = (uint64_t)blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        uint64_t r = by + j, c = bx + threadIdx.x;
        if (r < rows_in && c < cols_in) {
            uint32_t v = in[r * cols_in + c];
            if (conv == 1) v = Fp::from_canonical(v).v;
            else if (conv == 2) v = Fp::raw(v).to_canonical();
            tile[j][threadIdx.x] = v;
        }
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        uint64_t c = bx + j, r = by + threadIdx.x;
        if (r < rows_in && c < cols_in) out[c * rows_in + r] = tile[threadIdx.x][j];
    }
}
__global__ void transpose_tall_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t rows_in,
                                      uint64_t cols_in, int conv) {
    __shared__ uint32_t tile[32][33];
    uint64_t by = (uint64_t)blockIdx.x * 32, bx = (uint64_t)blockIdx.y * 32;  // rows on grid.x
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        uint64_t r = by + j, c = bx + threadIdx.x;
        if (r < rows_in && c < cols_in) {
            uint32_t v = in[r * cols_in + c];
            if (conv == 1) v = Fp::from_canonical(v).v;
            else if (conv == 2) v = Fp::raw(v).to_canonical();
            tile[j][threadIdx.x] = v;
        }
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        uint64_t c = bx + j, r = by + threadIdx.x;
        if (r < rows_in && c < cols_in) out[c * rows_in + r] = tile[threadIdx.x][j];
    }
}
// grid.y is limited to 65535 blocks, so the longer dimension goes on grid.x
static void launch_transpose(b2s_engine* e, const uint32_t* in, uint32_t* out, uint64_t rows_in, uint64_t cols_in,
                             int conv) {
    if (!rows_in || !cols_in) return;
    uint64_t rb = (rows_in + 31) / 32, cb = (cols_in + 31) / 32;
    if (rb <= 65535) {
        transpose_kernel<<<dim3((unsigned)cb, (unsigned)rb), dim3(32, 8), 0, e->stream>>>(in, out, rows_in, cols_in, conv);
    } else {
        B2_REQUIRE(cb <= 65535, B2S_ERR_SHAPE, "transpose: both dimensions exceed 2M");
        transpose_tall_kernel<<<dim3((unsigned)rb, (unsigned)cb), dim3(32, 8), 0, e->stream>>>(in, out, rows_in, cols_in, conv);
    }
    B2_CHECK_LAUNCH();
    e->launches[3]++;
}

void transpose_to_colmajor(b2s_engine* e, const uint32_t* d_rowmajor, uint32_t* d_colmajor, uint64_t height,
                           uint64_t width, bool to_monty) {
    // row-major (height x width)  ->  column-major = row-major (width x height)
    launch_transpose(e, d_rowmajor, d_colmajor, height, width, to_monty ? 1 : 0);
}
void transpose_to_rowmajor(b2s_engine* e, const uint32_t* d_colmajor, uint32_t* d_rowmajor, uint64_t height,
                           uint64_t width, bool to_canonical) {
    launch_transpose(e, d_colmajor, d_rowmajor, width, height, to_canonical ? 2 : 0);
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__global__ void fill_kernel(uint32_t* d, uint64_t height, uint64_t width, uint64_t seed) {
    uint64_t n = height * width;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        uint64_t c = i / height, r = i % height;
        uint32_t canon = (uint32_t)(splitmix64(seed ^ ((c << 32) | r)) % P);
        d[i] = Fp::from_canonical(canon).v;
    }
}
void fill_synthetic(b2s_engine* e, uint32_t* d, uint64_t height, uint64_t width, uint64_t seed) {
    if (!height || !width) return;
    fill_kernel<<<e->num_sms * 16, 256, 0, e->stream>>>(d, height, width, seed);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
}

// Satisfying trace of the authored chain AIR (openvm_b200/airs.py chain_air / chain_trace(hashed=True)), one thread per row:
// x_0 = pv0 + r, x_1 = hash(r, seed), x_{j+2} = x_j x_{j+1} + x_j for j + 2 <= W - 2, x_{W-1} = x_0 x_1 x_2. Stands in for a
// device tracegen in bench.py and the tests (the reference's tracegen is out of scope): a segment's 34 traces in 34 launches.
constexpr uint64_t CHAIN_LCG_A = 6364136223846793005ull, CHAIN_LCG_C = 1442695040888963407ull;
__host__ __device__ inline uint32_t chain_x1(uint64_t r, uint64_t seed) {
    const uint64_t x = (r + (seed * 1000003ull) % ((uint64_t)1 << 62)) * CHAIN_LCG_A + CHAIN_LCG_C;
    return (uint32_t)(((x >> 20) & (((uint64_t)1 << 40) - 1)) % P);
}
__global__ void chain_fill_kernel(uint32_t* d, uint64_t height, uint32_t width, uint32_t pv0, uint64_t seed) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= height) return;
    Fp a = Fp::from_canonical((uint32_t)((pv0 + r) % P)), b = Fp::from_canonical(chain_x1(r, seed));
    const Fp x0 = a, x1 = b;
    Fp x2 = Fp::zero();
    d[r] = a.v;
    d[height + r] = b.v;
    for (uint32_t j = 2; j + 1 < width; j++) {
        const Fp c = a * b + a;
        if (j == 2) x2 = c;
        d[(uint64_t)j * height + r] = c.v;
        a = b;
        b = c;
    }
    d[(uint64_t)(width - 1) * height + r] = (x0 * x1 * x2).v;
}
void fill_chain(b2s_engine* e, uint32_t* d, uint64_t height, uint64_t width, uint64_t seed, uint32_t pv[2]) {
    B2_REQUIRE(width >= 5 && width < ((uint64_t)1 << 31) && height >= 1, B2S_ERR_SHAPE, "fill_chain: width >= 5 and height >= 1");
    const uint32_t pv0 = (uint32_t)(((seed * CHAIN_LCG_A + CHAIN_LCG_C) >> 24) % P);
    chain_fill_kernel<<<(unsigned)((height + 127) / 128), 128, 0, e->stream>>>(d, height, (uint32_t)width, pv0, seed);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
    pv[0] = pv0;
    pv[1] = chain_x1(height - 1, seed);
}

// ---- Merkle opening gather ---------------------------------------------------------------------------
__global__ void gather_kernel(const OpenDesc* descs, uint32_t ndesc, uint32_t* out) {
    uint32_t di = blockIdx.x;
    if (di >= ndesc) return;
    OpenDesc d = descs[di];
    const bool raw = (d.count & OPEN_RAW) != 0;
    const uint32_t count = d.count & ~OPEN_RAW;
    for (uint32_t c = threadIdx.x; c < count; c += blockDim.x) {
        const uint32_t v = d.base[(uint64_t)c * d.stride + d.row];
        out[d.out_off + c] = raw ? v : Fp::raw(v).to_canonical();
    }
}
// every descriptor copies `count` strided words (a matrix row or a digest) to out[out_off..], canonical form;
// one launch and one device-to-host copy for the whole list
void gather_rows(b2s_engine* e, const std::vector<OpenDesc>& descs, size_t total_words, uint32_t* h_out) {
    if (descs.empty()) return;
    // both directions go through the engine's pinned buffer: a proof's query phase moves ~2 MB of descriptors down and
    // ~1.4 MB of rows and paths up, and pageable copies of that size cost several times the kernel
    const size_t db = descs.size() * sizeof(OpenDesc), ob = total_words * 4;
    const size_t db_pad = (db + 255) & ~(size_t)255;
    uint8_t* pin = (uint8_t*)e->host_pinned(db_pad + ob);
    memcpy(pin, descs.data(), db);
    OpenDesc* dd = (OpenDesc*)e->dmalloc(db);
    uint32_t* dout = (uint32_t*)e->dmalloc(ob);
    B2_CUDA(cudaMemcpyAsync(dd, pin, db, cudaMemcpyHostToDevice, e->stream));
    gather_kernel<<<(unsigned)descs.size(), 128, 0, e->stream>>>(dd, (uint32_t)descs.size(), dout);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
    B2_CUDA(cudaMemcpyAsync(pin + db_pad, dout, ob, cudaMemcpyDeviceToHost, e->stream));
    B2_CUDA(cudaStreamSynchronize(e->stream));
    memcpy(h_out, pin + db_pad, ob);
    e->dfree(dd, db);
    e->dfree(dout, ob);
}
void gather_open(b2s_engine* e, const b2s_pcs_data* d, uint64_t index, uint32_t* h_rows, uint32_t* h_proof) {
    B2_REQUIRE(index < ((uint64_t)1 << d->log_max_height), B2S_ERR_BAD_ARG, "open: index out of range");
    std::vector<OpenDesc> descs;
    uint32_t off = 0;
    for (auto* m : d->mats) {
        unsigned lh = log2_exact(m->height);
        descs.push_back(OpenDesc{m->d, m->height, index >> (d->log_max_height - lh), (uint32_t)m->width, off});
        off += (uint32_t)m->width;
    }
    uint32_t rows_words = off;
    for (unsigned l = 0; l < d->log_max_height; l++) {
        descs.push_back(OpenDesc{d->layers[l], d->layer_nodes[l], (index >> l) ^ 1, digest_words_desc(e), off});
        off += 8;
    }
    std::vector<uint32_t> tmp(off);
    gather_rows(e, descs, off, tmp.data());
    if (h_rows) memcpy(h_rows, tmp.data(), (size_t)rows_words * 4);
    if (h_proof) memcpy(h_proof, tmp.data() + rows_words, (size_t)(off - rows_words) * 4);
}

void download_layer(b2s_engine* e, const b2s_pcs_data* d, uint32_t level, uint32_t* h_out) {
    B2_REQUIRE(level < d->layers.size(), B2S_ERR_BAD_ARG, "layer: level out of range");
    uint64_t n = d->layer_nodes[level];
    // planes (8 x n) -> node-major (n x 8), canonical
    uint32_t* tmp = (uint32_t*)e->dmalloc(n * 32);
    transpose_to_rowmajor(e, d->layers[level], tmp, n, 8, e->hash_kind == 0);
    B2_CUDA(cudaMemcpyAsync(h_out, tmp, n * 32, cudaMemcpyDeviceToHost, e->stream));
    B2_CUDA(cudaStreamSynchronize(e->stream));
    e->dfree(tmp, n * 32);
}

// ---- integer-pipe micro-benchmarks -------------------------------------------------------------------
template <int ILP>
__global__ void modmul_bench_kernel(uint32_t* sink, uint32_t iters, uint32_t seed) {
    Fp a[ILP], b = Fp::raw(seed % P);
#pragma unroll
    for (int k = 0; k < ILP; k++) a[k] = Fp::raw((threadIdx.x * 2654435761u + k * 40503u + blockIdx.x) % P);
    // the cheapest product form this part has: signed Montgomery squarings (IMAD.WIDE + IMAD + IMAD.HI + subtract, no
    // correction into [0, p)), the form of the Poseidon2 S-box chains; 3.66 T/s against 3.1 T/s for the unsigned a * b form
    // (profiles/r1g_fp64_pipe.md). This is THE modular-product peak every int_pipe fraction in bench.py is quoted against.
    int32_t x[ILP];
#pragma unroll
    for (int k = 0; k < ILP; k++) x[k] = (int32_t)(a[k] * b).v - (int32_t)(P >> 1);
    for (uint32_t i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < ILP; k++) x[k] = smont(x[k], x[k]);
    }
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < ILP; k++) acc ^= (uint32_t)x[k];
    if (acc == 0x7FFFFFFFu) sink[0] = acc;  // |x| < p: never true; keeps the loop alive
}
double bench_modmul(b2s_engine* e, uint64_t threads, uint32_t iters, int ilp) {
    uint32_t* sink = (uint32_t*)e->dmalloc(4);
    cudaEvent_t a = e->get_event(), b = e->get_event();
    unsigned blocks = (unsigned)((threads + 255) / 256);
    B2_CUDA(cudaEventRecord(a, e->stream));
    switch (ilp) {
        case 1: modmul_bench_kernel<1><<<blocks, 256, 0, e->stream>>>(sink, iters, 12345u); break;
        case 2: modmul_bench_kernel<2><<<blocks, 256, 0, e->stream>>>(sink, iters, 12345u); break;
        case 4: modmul_bench_kernel<4><<<blocks, 256, 0, e->stream>>>(sink, iters, 12345u); break;
        default: modmul_bench_kernel<8><<<blocks, 256, 0, e->stream>>>(sink, iters, 12345u); break;
    }
    B2_CHECK_LAUNCH();
    B2_CUDA(cudaEventRecord(b, e->stream));
    B2_CUDA(cudaEventSynchronize(b));
    float ms = 0;
    B2_CUDA(cudaEventElapsedTime(&ms, a, b));
    e->event_pool.push_back(a);
    e->event_pool.push_back(b);
    e->dfree(sink, 4);
    e->launches[3]++;
    return ms;
}

__global__ void __launch_bounds__(128) poseidon2_bench_kernel(uint32_t* sink, uint32_t iters) {
    Fp s[16];
#pragma unroll
    for (int k = 0; k < 16; k++) s[k] = Fp::raw((threadIdx.x * 2654435761u + k * 40503u + blockIdx.x) % P);
    for (uint32_t i = 0; i < iters; i++) poseidon2_permute(s);
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) acc ^= s[k].v;
    if (acc == 0xFFFFFFFFu) sink[0] = acc;
}
double bench_poseidon2(b2s_engine* e, uint64_t threads, uint32_t iters) {
    uint32_t* sink = (uint32_t*)e->dmalloc(4);
    cudaEvent_t a = e->get_event(), b = e->get_event();
    unsigned blocks = (unsigned)((threads + 127) / 128);
    B2_CUDA(cudaEventRecord(a, e->stream));
    poseidon2_bench_kernel<<<blocks, 128, 0, e->stream>>>(sink, iters);
    B2_CHECK_LAUNCH();
    B2_CUDA(cudaEventRecord(b, e->stream));
    B2_CUDA(cudaEventSynchronize(b));
    float ms = 0;
    B2_CUDA(cudaEventElapsedTime(&ms, a, b));
    e->event_pool.push_back(a);
    e->event_pool.push_back(b);
    e->dfree(sink, 4);
    e->launches[3]++;
    e->permutations += (uint64_t)blocks * 128 * iters;
    return ms;
}

}  // namespace b2
