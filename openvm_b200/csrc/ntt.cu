// BabyBear coset low-degree extension: inverse NTT -> forward NTT on each coset, rows bit-reversed.
//
// Replaces (reference, external): the per-matrix part of Plonky3 `TwoAdicFriPcs::commit`
// (p3-fri / p3-dft `Radix2DitParallel::coset_lde_batch`), reached from
// crates/vm/src/arch/vm.rs:735 and crates/vm/src/system/program/trace.rs:58-63. Output convention pinned
// in-tree: leaf i of an LDE of height 2^h is the evaluation at GENERATOR * w_h^{bitrev_h(i)}
// (extensions/native/recursion/src/fri/two_adic_pcs.rs:178-186; domain: fri/domain.rs:142-154).
//
// Design (B200). On this part the transform is bound by the integer-multiply pipe and the dispatch slots its 64-bit
// forms hold, not by HBM (profiles/r2f_ncu_ntt_summary.md: math_pipe_throttle is the top stall of every pass, DRAM 37-50 %),
// so the kernels are built to execute exactly one product per butterfly, in its cheapest form, and nothing else:
//   * every stage is a Cooley-Tukey butterfly (a, b) -> (a + z*b, a - z*b), natural-order input, bit-reversed
//     output, with the stage twiddle AND the coset shift merged into one table entry z = Z[2^t + q]
//     (t = depth, q = sub-problem): no inter-digit twiddle products, no separate coset-scaling product;
//     the only extra product is the 1/N factor, applied once per coefficient;
//   * the product is Shoup's: the table holds (z, floor(z 2^32 / p)) and z*b mod p = b*z - mulhi(b, z')*p -- one IMAD.HI
//     and two 32-bit IMADs (4 + 2.1 + 2.1 issue cycles) instead of the Montgomery product's IMAD.WIDE + IMAD + IMAD.HI
//     (4.6 + 2.1 + 4); 2^24 x 256 LDE 97.4 -> 88.8 ms;
//   * tables: ZI (inverse, N pairs) and ZF (forward, one N-pair table per coset), built on the device once per
//     (log N, log_blowup, input shift) and cached in the engine; the 2^D - 1 pairs a tile needs are copied to shared memory
//     with the tile, so the butterfly loops contain no global loads;
//   * passes: [strided]* -> contiguous for the inverse (coefficients land bit-reversed), then
//     gather -> [strided]* -> contiguous for the forward transform of each coset. The gather pass reads the
//     bit-reversed coefficients as contiguous runs and writes natural slots, so no separate permutation pass
//     exists and the final output is in place in bit-reversed order. Coset j of the 2^b*N-point domain is
//     output block bitrev_b(j).
//   * strided / gather passes: 2^D x 32 tile in shared memory, a warp lane owns one of 32 adjacent positions,
//     radix-16 register blocks, twiddle index uniform per warp; contiguous pass: one run of 2^D <= 8192 words
//     per tile (several runs per block for small D), 1-in-32 padded shared memory, radix-16 register blocks.
#include "common.cuh"

#include <mutex>

namespace b2 {

constexpr int K1_THREADS = 256;
constexpr int K1_MAX_D = 9;    // 2^9 x 32 words = 64 KB tile
constexpr int K2_MAX_D = 13;   // 8192-word run

// ---- tables ---------------------------------------------------------------------------------------------
struct TabParams {
    uint32_t s[28];  // s[t] = base^(N / 2^(t+1))        (Montgomery)
    uint32_t g[28];  // g[t] = generator of order 2^(t+1) (forward or inverse)
};

// One entry = (w, w') with w the twiddle as a CANONICAL value and w' = floor(w 2^32 / p) (Shoup's precomputed quotient):
// v w mod p is then  v*w - mulhi(v, w')*p  in [0, 2p) -- one 4-cycle IMAD.HI and two 2-cycle IMADs instead of the
// Montgomery product's IMAD.WIDE + IMAD + IMAD.HI (4.6 + 2.1 + 4 cycles, profiles/r1j_dispatch_probe.md; the 64-bit forms
// hold the sub-partition's dispatch, and the r2f capture shows math_pipe_throttle as the top stall of every pass). The data
// words keep their Montgomery form: the map is linear.
__global__ void build_table_kernel(uint2* z, uint32_t n, TabParams p) {
    uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    if (idx == 0) { z[0] = make_uint2(0, 0); return; }
    const uint32_t t = 31 - __clz(idx);
    const uint32_t q = idx - (1u << t);
    const uint32_t e = t ? (__brev(q) >> (32 - t)) : 0u;
    const uint32_t w = (Fp::raw(p.s[t]) * fp_pow(Fp::raw(p.g[t]), e)).to_canonical();
    z[idx] = make_uint2(w, (uint32_t)(((uint64_t)w << 32) / P));
}

struct TwiddleSet {  // inverse table for one log size
    unsigned logn;
    uint2* zi;
};
struct ScaleSet {    // forward tables (one per coset) for (logn, log_blowup, shift_in)
    unsigned logn, log_blowup;
    uint2* zf;       // [2^b][N]
    uint32_t ninv;   // 1/N, Montgomery
};

static void fill_table(b2s_engine* e, uint2* z, unsigned logn, Fp base, bool inverse) {
    TabParams p{};
    for (unsigned t = 0; t < logn; t++) {
        Fp g = two_adic_generator(t + 1);
        p.g[t] = (inverse ? fp_inv(g) : g).v;
        p.s[t] = fp_pow(base, (uint64_t)1 << (logn - t - 1)).v;
    }
    const uint32_t n = 1u << logn;
    build_table_kernel<<<(n + 255) / 256, 256, 0, e->stream>>>(z, n, p);
    B2_CHECK_LAUNCH();
    e->launches[3]++;
}

static TwiddleSet* get_twiddles(b2s_engine* e, unsigned logn) {
    auto it = e->twiddles.find(logn);
    if (it != e->twiddles.end()) return it->second;
    TwiddleSet* t = new TwiddleSet();
    t->logn = logn;
    t->zi = (uint2*)e->dmalloc((size_t)8 << logn);
    fill_table(e, t->zi, logn, Fp::one(), true);
    // built once per engine: complete before any other stream of the engine (side-stream LDEs of one commit) can find it cached
    B2_CUDA(cudaStreamSynchronize(e->stream));
    e->twiddles[logn] = t;
    return t;
}

void free_twiddles(b2s_engine* e) {
    for (auto& kv : e->twiddles) {
        e->dfree(kv.second->zi, (size_t)8 << kv.second->logn);
        delete kv.second;
    }
    e->twiddles.clear();
}

static ScaleSet* get_scales(b2s_engine* e, unsigned logn, unsigned log_blowup, Fp shift_in) {
    uint64_t key = ((uint64_t)logn << 40) | ((uint64_t)log_blowup << 32) | shift_in.v;
    auto it = e->scales.find(key);
    if (it != e->scales.end()) return it->second;
    ScaleSet* s = new ScaleSet();
    s->logn = logn;
    s->log_blowup = log_blowup;
    const uint32_t B = 1u << log_blowup;
    s->zf = (uint2*)e->dmalloc(((size_t)8 << logn) * B);
    Fp base = Fp::raw(M_GENERATOR) * fp_inv(shift_in);
    Fp wm = two_adic_generator(logn + log_blowup);
    for (uint32_t j = 0; j < B; j++) {
        fill_table(e, s->zf + ((size_t)j << logn), logn, base, false);
        base *= wm;
    }
    s->ninv = fp_inv(Fp::from_canonical((uint32_t)(((uint64_t)1 << logn) % P))).v;
    B2_CUDA(cudaStreamSynchronize(e->stream));   // as for the inverse table
    e->scales[key] = s;
    return s;
}

void free_scales(b2s_engine* e) {
    for (auto& kv : e->scales) {
        ScaleSet* s = kv.second;
        e->dfree(s->zf, ((size_t)8 << s->logn) << s->log_blowup);
        delete s;
    }
    e->scales.clear();
}

// ---- register butterfly block -----------------------------------------------------------------------------
// 2^E consecutive table words starting at z (aligned to 4 * 2^E bytes)
// 2^E consecutive words starting at z (aligned to 4 * 2^E bytes), from the tile's twiddles in shared memory
template <int E>
B2_D void load_tw(const uint32_t* z, uint32_t (&w)[1 << E]) {
    if constexpr (E == 0) {
        w[0] = z[0];
    } else if constexpr (E == 1) {
        uint2 a = *reinterpret_cast<const uint2*>(z);
        w[0] = a.x; w[1] = a.y;
    } else if constexpr (E == 2) {
        uint4 a = *reinterpret_cast<const uint4*>(z);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
    } else {
        uint4 a = *reinterpret_cast<const uint4*>(z);
        uint4 b = *(reinterpret_cast<const uint4*>(z) + 1);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
        w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    }
}

// The 2^D - 1 table entries a tile's D depths use, copied next to the tile: tw[2^k + i] = Z[(root << k) + i] for depth k of the
// tile, i < 2^k, root = 2^t0 + (sub-problem of the tile); the Shoup companions go to tw[count + ...]. The butterfly loops read
// their twiddles from shared memory (index (1 << depth) + sub-problem inside the tile); read straight from the table they
// were L2-latency loads in the middle of the arithmetic (r1e: long_scoreboard 5.6 -> r2f: 1.3).
B2_D void stage_twiddles(uint32_t* tw, const uint2* __restrict__ z, uint32_t root, uint32_t count, uint32_t tid, uint32_t nthreads) {
    for (uint32_t i = tid; i < count; i += nthreads) {
        uint2 v = make_uint2(0, 0);
        if (i) {
            const uint32_t k = 31 - __clz(i);
            v = __ldg(z + ((size_t)root << k) + (i - (1u << k)));
        }
        tw[i] = v.x;
        tw[count + i] = v.y;
    }
}

// v * w mod p for a canonical twiddle w with its Shoup companion ws; v in [0, p) in any (linear) representation
B2_D Fp shoup_mul(Fp v, uint32_t w, uint32_t ws) {
    const uint32_t q = __umulhi(v.v, ws);
    const uint32_t r = v.v * w - q * P;       // exact value in [0, 2p), computed mod 2^32
    const uint32_t c = r - P;
    return Fp::raw(c < r ? c : r);
}

// tws = tw + TW_COUNT (the companions): TWC is the number of entries per plane (2^D)
template <int R, int E, int TWC>
B2_D void ct_stage(Fp (&v)[1 << R], const uint32_t* z, uint32_t idx0) {
    constexpr int half = 1 << (R - 1 - E);
    uint32_t w[1 << E], ws[1 << E];
    load_tw<E>(z + ((size_t)idx0 << E), w);
    load_tw<E>(z + TWC + ((size_t)idx0 << E), ws);
#pragma unroll
    for (int j = 0; j < (1 << E); j++) {
#pragma unroll
        for (int k = 0; k < half; k++) {
            const int ia = j * 2 * half + k, ib = ia + half;
            Fp t = shoup_mul(v[ib], w[j], ws[j]);
            Fp a = v[ia];
            v[ia] = a + t;
            v[ib] = a - t;
        }
    }
}

// R depths on 2^R values; depth e pairs indices differing in bit R-1-e; idx0 = 2^depth + sub-problem
template <int R, int TWC>
B2_D void ct_block(Fp (&v)[1 << R], const uint32_t* z, uint32_t idx0) {
    ct_stage<R, 0, TWC>(v, z, idx0);
    if constexpr (R >= 2) ct_stage<R, 1, TWC>(v, z, idx0);
    if constexpr (R >= 3) ct_stage<R, 2, TWC>(v, z, idx0);
    if constexpr (R >= 4) ct_stage<R, 3, TWC>(v, z, idx0);
}

template <int R>
B2_D constexpr int brev_c(int m) {
    int r = 0;
    for (int i = 0; i < R; i++) r |= ((m >> i) & 1) << (R - 1 - i);
    return r;
}

constexpr int first_radix(int D) { return D == 0 ? 0 : ((D & 3) ? (D & 3) : 4); }

// ---- strided / gather pass ----------------------------------------------------------------------------------
struct PassParams {
    const uint32_t* in;
    uint32_t* out;
    uint64_t in_col_stride, out_col_stride;
    const uint2* z;           // table of (w, w') pairs; coset c uses z + c * 2^log_len
    uint32_t width;           // columns covered by the launch (contiguous pass: for the run -> column map)
    uint32_t log_len;         // log2 of the transform length n
    uint32_t t0;              // global depth of the first stage of this pass
    uint32_t log_blowup;      // coset c writes (and, when in_block, reads) block bitrev_b(c) of 2^log_len words
    uint32_t coset;           // 1: blockIdx.z is the coset index
    uint32_t in_block;        // 1: the input is offset by the coset block as well (in-place passes on the LDE)
    uint32_t scale;           // != 0: multiply loaded words by this (1/N; first forward pass)
    uint32_t reverse_load;    // contiguous pass with D == log_len: word p of the run is stored at slot bitrev_D(p)
    // Row-sharded output over peer memory (last pass of a sharded commitment, openvm_b200/sharded.py): LDE row `row` of
    // launch column `col` is stored at peer_out[row >> peer_log_rows] + (peer_col0 + col) * 2^peer_log_rows + (row mod
    // 2^peer_log_rows) -- the buffer of the rank that will hash that row, written over NVLink by this kernel's own stores.
    uint32_t peer_n;          // 0: plain output
    uint32_t peer_log_rows;
    uint64_t peer_col0;
    uint32_t* peer_out[B2_MAX_PEERS];
};

template <int D, int U, bool GATHER>
struct K1Rounds {
    static B2_D void run(uint32_t* x, const uint32_t* z, uint32_t sub0) {
        constexpr int R = (U == 0) ? first_radix(D) : 4;
        constexpr int PITCH = GATHER ? 33 : 32;
        constexpr uint32_t NBLK = 1u << (D - R);
        const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        for (uint32_t blk = warp; blk < NBLK; blk += K1_THREADS / 32) {
            Fp v[1 << R];
            uint32_t row0, rs, idx0;
            if constexpr (GATHER) {
                // depths U..U+R-1 pair slots at distance 2^U, 2^(U+1), ...: slot m of the block is register bitrev_R(m)
                const uint32_t lo = blk & ((1u << U) - 1), hi = blk >> U;
                row0 = (hi << (R + U)) + lo;
                rs = 1u << U;
                idx0 = (1u << U) + (U ? (__brev(lo) >> (32 - U)) : 0u);
#pragma unroll
                for (int m = 0; m < (1 << R); m++) v[brev_c<R>(m)] = Fp::raw(x[(row0 + m * rs) * PITCH + lane]);
            } else {
                constexpr int LS = D - U - R;  // log2 of the row stride inside a block
                const uint32_t lo = blk & ((1u << LS) - 1), hi = blk >> LS;
                row0 = (hi << (R + LS)) + lo;
                rs = 1u << LS;
                idx0 = (sub0 << U) + hi;
#pragma unroll
                for (int m = 0; m < (1 << R); m++) v[m] = Fp::raw(x[(row0 + m * rs) * PITCH + lane]);
            }
            ct_block<R, (1 << D)>(v, z, idx0);
            if constexpr (GATHER) {
#pragma unroll
                for (int m = 0; m < (1 << R); m++) x[(row0 + m * rs) * PITCH + lane] = v[brev_c<R>(m)].v;
            } else {
#pragma unroll
                for (int m = 0; m < (1 << R); m++) x[(row0 + m * rs) * PITCH + lane] = v[m].v;
            }
        }
        __syncthreads();
        K1Rounds<D, U + R, GATHER>::run(x, z, sub0);
    }
};
template <int D, bool GATHER>
struct K1Rounds<D, D, GATHER> {
    static B2_D void run(uint32_t*, const uint32_t*, uint32_t) {}
};

// Tile: rows r in [0, 2^D) at stride 2^(n - t0 - D) inside sub-problem `top` of depth t0, 32 adjacent positions.
// GATHER (first forward pass, t0 = 0): the input holds coefficient i at slot bitrev_n(i); row r / lane t of the tile
// is read from run bitrev_{n-D}(32*chunk + t), word r, and the result of slot r goes to natural row bitrev_D(r).
template <int D, bool GATHER>
__global__ void __launch_bounds__(K1_THREADS) ntt_strided_kernel(PassParams p) {
    extern __shared__ uint32_t x[];
    constexpr int PITCH = GATHER ? 33 : 32;
    const uint32_t n = p.log_len;
    const uint32_t low_bits = n - p.t0 - D;            // >= 5
    const uint32_t chunk = blockIdx.x & ((1u << (low_bits - 5)) - 1);
    const uint32_t top = blockIdx.x >> (low_bits - 5);
    const uint32_t col = blockIdx.y;
    const uint32_t cz = p.coset ? blockIdx.z : 0;
    const uint64_t blk_off = p.coset && p.log_blowup ? ((uint64_t)(__brev(cz) >> (32 - p.log_blowup)) << n) : 0;
    const uint2* __restrict__ z = p.z + ((size_t)cz << n);
    const uint32_t tid = threadIdx.x;
    const uint32_t* in = p.in + (uint64_t)col * p.in_col_stride + (p.in_block ? blk_off : 0);
    uint32_t* out = p.out + (uint64_t)col * p.out_col_stride + blk_off;

    // Staging moves 16 bytes per thread and iteration (LDG.128 / STG.128): the index arithmetic, which is a quarter of
    // all instructions of a pass when done per word (profiles/r1c), is paid once per four words.
    if constexpr (GATHER) {
        const Fp sc = Fp::raw(p.scale);
        if constexpr (D >= 2) {
            // a warp covers 2^A 16-byte pieces of 2^(5-A) runs: full 128-byte lines from global memory, and the four
            // transposed stores land in banks 4a + b (pitch 33), all different for A = 3
            constexpr int A = (D - 2) < 3 ? (D - 2) : 3;
            constexpr uint32_t TOTAL = 8u << D;
            constexpr int BATCH = TOTAL / K1_THREADS >= 8 ? 8 : (TOTAL / K1_THREADS ? TOTAL / K1_THREADS : 1);
            for (uint32_t e0 = 0; e0 < TOTAL; e0 += BATCH * K1_THREADS) {
                uint4 buf[BATCH];
#pragma unroll
                for (int k = 0; k < BATCH; k++) {
                    const uint32_t e = e0 + k * K1_THREADS + tid;
                    const uint32_t a = e & ((1u << A) - 1), b = (e >> A) & ((1u << (5 - A)) - 1);
                    const uint32_t c = (e >> 5) & ((1u << (D - 2 - A)) - 1), d = e >> (D + 3 - A);
                    const uint32_t r = ((c << A) | a) << 2, t = (d << (5 - A)) | b;
                    const uint32_t ilo = (chunk << 5) + t;
                    const uint64_t run = (uint64_t)(__brev(ilo) >> (32 - low_bits)) << D;
                    if (e < TOTAL) buf[k] = *reinterpret_cast<const uint4*>(in + run + r);
                }
#pragma unroll
                for (int k = 0; k < BATCH; k++) {
                    const uint32_t e = e0 + k * K1_THREADS + tid;
                    const uint32_t a = e & ((1u << A) - 1), b = (e >> A) & ((1u << (5 - A)) - 1);
                    const uint32_t c = (e >> 5) & ((1u << (D - 2 - A)) - 1), d = e >> (D + 3 - A);
                    const uint32_t r = ((c << A) | a) << 2, t = (d << (5 - A)) | b;
                    if (e < TOTAL) {
                        uint32_t* xd = x + r * PITCH + t;
                        xd[0] = (Fp::raw(buf[k].x) * sc).v;
                        xd[PITCH] = (Fp::raw(buf[k].y) * sc).v;
                        xd[2 * PITCH] = (Fp::raw(buf[k].z) * sc).v;
                        xd[3 * PITCH] = (Fp::raw(buf[k].w) * sc).v;
                    }
                }
            }
        } else {
            for (uint32_t e = tid; e < (32u << D); e += K1_THREADS) {
                const uint32_t r = e & ((1u << D) - 1), t = e >> D;
                const uint32_t ilo = (chunk << 5) + t;
                const uint64_t run = (uint64_t)(__brev(ilo) >> (32 - low_bits)) << D;
                x[r * PITCH + t] = (Fp::raw(in[run + r]) * sc).v;
            }
        }
    } else {
        // all loads of a batch are issued before the first store (up to 8 x 16 bytes in flight per thread): the loop
        // as written per element was one dependent LDG -> STS round trip per iteration (r1e: long_scoreboard 5.6)
        const uint64_t base = ((uint64_t)top << (n - p.t0)) + ((uint64_t)chunk << 5);
        constexpr uint32_t TOTAL = 8u << D;
        constexpr int BATCH = TOTAL / K1_THREADS >= 8 ? 8 : (TOTAL / K1_THREADS ? TOTAL / K1_THREADS : 1);
        for (uint32_t e0 = 0; e0 < TOTAL; e0 += BATCH * K1_THREADS) {
            uint4 buf[BATCH];
#pragma unroll
            for (int k = 0; k < BATCH; k++) {
                const uint32_t e = e0 + k * K1_THREADS + tid;
                const uint32_t q = (e & 7) << 2, r = e >> 3;
                if (e < TOTAL) buf[k] = *reinterpret_cast<const uint4*>(in + base + ((uint64_t)r << low_bits) + q);
            }
#pragma unroll
            for (int k = 0; k < BATCH; k++) {
                const uint32_t e = e0 + k * K1_THREADS + tid;
                const uint32_t q = (e & 7) << 2, r = e >> 3;
                if (e < TOTAL) *reinterpret_cast<uint4*>(x + r * PITCH + q) = buf[k];
            }
        }
    }
    uint32_t* tw = x + ((size_t)PITCH << D);
    stage_twiddles(tw, z, (1u << p.t0) + top, 1u << D, tid, K1_THREADS);
    __syncthreads();
    K1Rounds<D, 0, GATHER>::run(x, tw, 1u);
    if constexpr (GATHER) {
        const uint64_t base = (uint64_t)chunk << 5;
        for (uint32_t e = tid; e < (8u << D); e += K1_THREADS) {
            const uint32_t q = (e & 7) << 2, r = e >> 3;
            const uint32_t row = D ? (__brev(r) >> (32 - D)) : 0u;
            const uint32_t* xs = x + r * PITCH + q;
            *reinterpret_cast<uint4*>(out + base + ((uint64_t)row << low_bits) + q) = make_uint4(xs[0], xs[1], xs[2], xs[3]);
        }
    } else {
        const uint64_t base = ((uint64_t)top << (n - p.t0)) + ((uint64_t)chunk << 5);
        for (uint32_t e = tid; e < (8u << D); e += K1_THREADS) {
            const uint32_t q = (e & 7) << 2, r = e >> 3;
            *reinterpret_cast<uint4*>(out + base + ((uint64_t)r << low_bits) + q) = *reinterpret_cast<const uint4*>(x + r * PITCH + q);
        }
    }
}

// ---- strided pass, asynchronous variant (experiment, B2S_NTT_ASYNC=1) ---------------------------------------------------
// Persistent blocks, two tile buffers: the 16-byte pieces of the block's NEXT tile travel global -> shared with cp.async
// (LDGSTS: no registers, no LDG/STS pair) while the butterflies of the current tile run; its twiddle pair sits in two
// registers per thread meanwhile. This is the cp.async form of the TMA / mbarrier double buffer (a 2-D TMA box of 32 words x
// 2^D rows would move the same bytes): what it can gain is load latency hidden inside one block, what it costs is shared
// memory (2 tiles -> 3 blocks per SM instead of 5-6). Measured in profiles/r2zb_ntt_async.md.
B2_D void cp_async16(uint32_t* smem_dst, const uint32_t* gsrc) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gsrc));
}
B2_D void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
B2_D void cp_async_wait_all_but_one() { asm volatile("cp.async.wait_group 1;\n" ::: "memory"); }

template <int D>
__global__ void __launch_bounds__(K1_THREADS) ntt_strided_async_kernel(PassParams p, uint32_t log_tiles_x, uint32_t total_tiles) {
    extern __shared__ uint32_t smem[];
    constexpr uint32_t TILE = 32u << D, TWW = 2u << D, COUNT = 1u << D;
    constexpr uint32_t TOTAL = 8u << D;                                  // 16-byte pieces per tile
    constexpr int NB = TOTAL / K1_THREADS ? TOTAL / K1_THREADS : 1;
    constexpr int NT = (COUNT + K1_THREADS - 1) / K1_THREADS;            // twiddle pairs per thread
    const uint32_t n = p.log_len;
    const uint32_t low_bits = n - p.t0 - D;
    const uint32_t tid = threadIdx.x;
    struct Tile {
        uint32_t chunk, top;
        const uint32_t* in;
        uint32_t* out;
        const uint2* z;
    };
    auto decode = [&](uint32_t T) {
        Tile tl;
        const uint32_t xi = T & ((1u << log_tiles_x) - 1), rest = T >> log_tiles_x;
        const uint32_t col = rest % p.width, cz = rest / p.width;
        tl.chunk = xi & ((1u << (low_bits - 5)) - 1);
        tl.top = xi >> (low_bits - 5);
        const uint64_t blk_off = p.coset && p.log_blowup ? ((uint64_t)(__brev(cz) >> (32 - p.log_blowup)) << n) : 0;
        tl.z = p.z + ((size_t)cz << n);
        tl.in = p.in + (uint64_t)col * p.in_col_stride + (p.in_block ? blk_off : 0);
        tl.out = p.out + (uint64_t)col * p.out_col_stride + blk_off;
        return tl;
    };
    uint2 twr[NT];
    auto issue = [&](const Tile& tl, uint32_t* xb) {
        const uint64_t base = ((uint64_t)tl.top << (n - p.t0)) + ((uint64_t)tl.chunk << 5);
#pragma unroll
        for (int k = 0; k < NB; k++) {
            const uint32_t e = k * K1_THREADS + tid;
            const uint32_t q = (e & 7) << 2, r = e >> 3;
            if (e < TOTAL) cp_async16(xb + r * 32 + q, tl.in + base + ((uint64_t)r << low_bits) + q);
        }
        const uint32_t root = (1u << p.t0) + tl.top;
#pragma unroll
        for (int k = 0; k < NT; k++) {
            const uint32_t i = k * K1_THREADS + tid;
            twr[k] = make_uint2(0, 0);
            if (i && i < COUNT) {
                const uint32_t lv = 31 - __clz(i);
                twr[k] = __ldg(tl.z + ((size_t)root << lv) + (i - (1u << lv)));
            }
        }
    };
    uint32_t T = blockIdx.x, buf = 0;
    Tile cur{};
    if (T < total_tiles) {
        cur = decode(T);
        issue(cur, smem);
    }
    cp_async_commit();
    while (T < total_tiles) {
        uint32_t* xb = smem + buf * (TILE + TWW);
        uint32_t* tw = xb + TILE;
#pragma unroll
        for (int k = 0; k < NT; k++) {
            const uint32_t i = k * K1_THREADS + tid;
            if (i < COUNT) { tw[i] = twr[k].x; tw[COUNT + i] = twr[k].y; }
        }
        const uint32_t Tn = T + gridDim.x;
        Tile nxt{};
        if (Tn < total_tiles) {
            nxt = decode(Tn);
            issue(nxt, smem + (buf ^ 1) * (TILE + TWW));
        }
        cp_async_commit();            // possibly empty: the wait below then still covers the current tile's group
        cp_async_wait_all_but_one();
        __syncthreads();
        K1Rounds<D, 0, false>::run(xb, tw, 1u);
        const uint64_t base = ((uint64_t)cur.top << (n - p.t0)) + ((uint64_t)cur.chunk << 5);
        for (uint32_t e = tid; e < TOTAL; e += K1_THREADS) {
            const uint32_t q = (e & 7) << 2, r = e >> 3;
            *reinterpret_cast<uint4*>(cur.out + base + ((uint64_t)r << low_bits) + q) = *reinterpret_cast<const uint4*>(xb + r * 32 + q);
        }
        __syncthreads();
        T = Tn;
        cur = nxt;
        buf ^= 1;
    }
}

// ---- contiguous pass ------------------------------------------------------------------------------------------
template <int D>
struct K2Cfg {
    static constexpr int TPR = D >= 4 ? (1 << (D - 4)) : 1;          // threads per run
    static constexpr int THREADS = TPR > 256 ? TPR : 256;
    static constexpr int RPC = THREADS / TPR;                          // runs per block
    static constexpr int PITCH = (1 << D) + (D >= 5 ? (1 << (D - 5)) : 1);
};
B2_D uint32_t phys(uint32_t l) { return l + (l >> 5); }
template <int LS, int R>
B2_D constexpr int pad_off(int m) {
    return (m << LS) + (LS >= 5 ? (m << (LS >= 5 ? LS - 5 : 0)) : (LS + R >= 5 ? (m >> (LS >= 5 ? 0 : 5 - LS)) : 0));
}

template <int D, int U>
struct K2Rounds {
    static B2_D void run(uint32_t* xr, const uint32_t* z, uint32_t sub0, uint32_t tr, bool live) {
        constexpr int R = (U == 0) ? first_radix(D) : 4;
        constexpr int LS = D - U - R;
        constexpr uint32_t NBLK = 1u << (D - R);
        if (live) {
            for (uint32_t blk = tr; blk < NBLK; blk += K2Cfg<D>::TPR) {
                const uint32_t lo = blk & ((1u << LS) - 1);
                uint32_t hi = blk >> LS;
                // row stride 16: take lane bit 4 from hi bit 1 so that the 32 lanes fall into 32 different banks
                if (LS == 4 && U >= 2) hi = (hi & ~3u) | ((hi & 1u) << 1) | ((hi >> 1) & 1u);
                const uint32_t l0 = (hi << (R + LS)) + lo;
                // phys(l0 + (m << LS)) = phys(l0) + a compile-time offset: bits [LS, LS+R) of l0 are zero
                uint32_t* xb = xr + phys(l0);
                Fp v[1 << R];
#pragma unroll
                for (int m = 0; m < (1 << R); m++) v[m] = Fp::raw(xb[pad_off<LS, R>(m)]);
                ct_block<R, (1 << D)>(v, z, (sub0 << U) + hi);
#pragma unroll
                for (int m = 0; m < (1 << R); m++) xb[pad_off<LS, R>(m)] = v[m].v;
            }
        }
        __syncthreads();
        K2Rounds<D, U + R>::run(xr, z, sub0, tr, live);
    }
};
template <int D>
struct K2Rounds<D, D> {
    static B2_D void run(uint32_t*, const uint32_t*, uint32_t, uint32_t, bool) {}
};

// One run = 2^D consecutive words of a column = sub-problem `rho` of depth t0 = n - D. Runs are numbered across
// the columns of the launch, so tiny heights still fill a block.
template <int D>
__global__ void __launch_bounds__(K2Cfg<D>::THREADS) ntt_contig_kernel(PassParams p) {
    extern __shared__ uint32_t x[];
    using C = K2Cfg<D>;
    const uint32_t n = p.log_len;
    const uint32_t log_rpc = n - D;  // runs per column
    const uint64_t total_runs = (uint64_t)p.width << log_rpc;
    const uint32_t cz = p.coset ? blockIdx.z : 0;
    const uint64_t blk_off = p.coset && p.log_blowup ? ((uint64_t)(__brev(cz) >> (32 - p.log_blowup)) << n) : 0;
    const uint64_t in_off = p.in_block ? blk_off : 0;
    const uint2* __restrict__ z = p.z + ((size_t)cz << n);
    const uint32_t tid = threadIdx.x;
    // Runs are numbered column-fastest (g = rho * width + col): the runs of one block, and neighbouring blocks, work on
    // the same sub-problem rho of different columns and therefore read the same twiddles (L1 / L2 hits instead of one
    // pass over the 2^n-word table per column; r1e capture: the contiguous pass read 2x its data from DRAM).
    const uint64_t run0 = (uint64_t)blockIdx.x * C::RPC;
    const uint32_t rho0 = (uint32_t)(run0 / p.width), col0 = (uint32_t)(run0 % p.width);
    auto locate = [&](uint32_t rs, uint32_t& col, uint32_t& rho) {
        const uint32_t c = col0 + rs;
        if ((p.width & (p.width - 1)) == 0) {  // column groups are powers of two except a ragged last one
            col = c & (p.width - 1);
            rho = rho0 + (c >> (31 - __clz(p.width)));
        } else if (p.width >= (uint32_t)C::RPC) {
            const uint32_t wrap = c >= p.width ? 1u : 0u;
            col = c - wrap * p.width;
            rho = rho0 + wrap;
        } else {
            col = c % p.width;
            rho = rho0 + c / p.width;
        }
    };
    const Fp sc = Fp::raw(p.scale);

    // 16 bytes per thread and iteration when the run is at least 4 words (see ntt_strided_kernel)
    constexpr int V = D >= 2 ? 4 : 1;
    constexpr int LV = D >= 2 ? 2 : 0;
    constexpr uint32_t TOTAL = (uint32_t)C::RPC << (D - LV);
    constexpr int BATCH = TOTAL / C::THREADS >= 4 ? 4 : (TOTAL / C::THREADS ? TOTAL / C::THREADS : 1);
    for (uint32_t e0 = 0; e0 < TOTAL; e0 += BATCH * C::THREADS) {
        uint32_t v[BATCH][V];
#pragma unroll
        for (int k = 0; k < BATCH; k++) {  // all loads of the batch in flight before the first use
            const uint32_t e = e0 + k * C::THREADS + tid;
            const uint32_t rs = e >> (D - LV), w = (e & ((1u << (D - LV)) - 1)) << LV;
            uint32_t col, rho;
            locate(rs, col, rho);
            if (e < TOTAL && run0 + rs < total_runs) {
                const uint32_t* src = p.in + (uint64_t)col * p.in_col_stride + in_off + ((uint64_t)rho << D) + w;
                if constexpr (V == 4) {
                    const uint4 q = *reinterpret_cast<const uint4*>(src);
                    v[k][0] = q.x; v[k][1] = q.y; v[k][2] = q.z; v[k][3] = q.w;
                } else {
                    v[k][0] = src[0];
                }
            }
        }
#pragma unroll
        for (int k = 0; k < BATCH; k++) {
            const uint32_t e = e0 + k * C::THREADS + tid;
            const uint32_t rs = e >> (D - LV), w = (e & ((1u << (D - LV)) - 1)) << LV;
            if (e < TOTAL && run0 + rs < total_runs) {
                uint32_t* xr = x + rs * C::PITCH;
#pragma unroll
                for (int j = 0; j < V; j++) {
                    const uint32_t val = p.scale ? (Fp::raw(v[k][j]) * sc).v : v[k][j];
                    const uint32_t slot = (p.reverse_load && D) ? (__brev(w + j) >> (32 - D)) : w + j;
                    xr[phys(slot)] = val;
                }
            }
        }
    }
    // twiddles of every sub-problem rho this block touches (runs are numbered column-fastest: at most RPC of them, usually
    // RPC / width + 1), one 2^D-word set each
    uint32_t* tw = x + (size_t)C::RPC * C::PITCH;
    {
        uint32_t nrho = (col0 + C::RPC - 1) / p.width + 1;
        if (nrho > (uint32_t)C::RPC) nrho = C::RPC;
        const uint32_t max_rho = 1u << log_rpc;
        for (uint32_t j = 0; j < nrho && rho0 + j < max_rho; j++)
            stage_twiddles(tw + ((size_t)j << (D + 1)), z, (1u << p.t0) + rho0 + j, 1u << D, tid, C::THREADS);
    }
    __syncthreads();
    {
        const uint32_t rs = tid / C::TPR, tr = tid % C::TPR;
        const uint64_t g = run0 + rs;
        uint32_t col, rho;
        locate(rs, col, rho);
        K2Rounds<D, 0>::run(x + rs * C::PITCH, tw + ((size_t)(rho - rho0) << (D + 1)), 1u, tr, g < total_runs);
    }
    for (uint32_t e = tid; e < ((uint32_t)C::RPC << (D - LV)); e += C::THREADS) {
        const uint32_t rs = e >> (D - LV), w = (e & ((1u << (D - LV)) - 1)) << LV;
        const uint64_t g = run0 + rs;
        if (g >= total_runs) break;
        uint32_t col, rho;
        locate(rs, col, rho);
        uint32_t* dst;
        if (p.peer_n) {
            const uint64_t row = blk_off + ((uint64_t)rho << D) + w;   // 4 words never straddle two peers' row blocks
            dst = p.peer_out[row >> p.peer_log_rows] + ((p.peer_col0 + col) << p.peer_log_rows) +
                  (row & (((uint64_t)1 << p.peer_log_rows) - 1));
        } else {
            dst = p.out + (uint64_t)col * p.out_col_stride + blk_off + ((uint64_t)rho << D) + w;
        }
        const uint32_t* xs = x + rs * C::PITCH + phys(w);  // w % 4 == 0: the four words share one group of 32
        if constexpr (V == 4) *reinterpret_cast<uint4*>(dst) = make_uint4(xs[0], xs[1], xs[2], xs[3]);
        else dst[0] = xs[0];
    }
}

// height-1 matrices: the polynomial is the constant; every extended row equals the single input row
__global__ void replicate_kernel(const uint32_t* in, uint64_t in_col_stride, uint32_t* out, uint64_t out_col_stride,
                                 uint32_t width, uint32_t rows) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= width * rows) return;
    uint32_t c = i / rows, r = i % rows;
    out[(uint64_t)c * out_col_stride + r] = in[(uint64_t)c * in_col_stride];
}

// ---- host side --------------------------------------------------------------------------------------------------
static bool ntt_async_enabled() {
    static const bool v = [] { const char* s = getenv("B2S_NTT_ASYNC"); return s && s[0] == '1'; }();
    return v;
}
template <int D>
static void launch_strided_async_t(b2s_engine* e, const PassParams& p, dim3 grid) {
    constexpr size_t smem = 2 * ((size_t)(32u << D) + (2u << D)) * 4;
    static std::once_flag configured[64];
    static unsigned resident[64];
    cudaError_t err = cudaSuccess;
    std::call_once(configured[e->device & 63], [&] {
        err = cudaFuncSetAttribute(ntt_strided_async_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        int per_sm = 0;
        if (err == cudaSuccess) err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ntt_strided_async_kernel<D>, K1_THREADS, smem);
        resident[e->device & 63] = (unsigned)(e->num_sms * (per_sm < 1 ? 1 : per_sm));
    });
    B2_CUDA(err);
    const uint64_t total = (uint64_t)grid.x * grid.y * grid.z;
    B2_REQUIRE(total <= 0xFFFFFFFFull && (grid.x & (grid.x - 1)) == 0, B2S_ERR_SHAPE, "ntt: too many tiles in one launch");
    const unsigned blocks = (unsigned)std::min<uint64_t>(total, resident[e->device & 63]);
    ntt_strided_async_kernel<D><<<blocks, K1_THREADS, smem, e->stream>>>(p, log2_exact(grid.x), (uint32_t)total);
    B2_CHECK_LAUNCH();
    e->launches[0]++;
}
template <int D, bool GATHER>
static void launch_strided_t(b2s_engine* e, const PassParams& p, dim3 grid) {
    if constexpr (!GATHER && D >= 5 && D <= 8) {
        if (ntt_async_enabled()) return launch_strided_async_t<D>(e, p, grid);
    }
    constexpr size_t smem = ((size_t)1 << D) * ((GATHER ? 33 : 32) + 2) * 4;   // tile + its twiddles (w and w')
    if (smem > 48 * 1024) {
        // once per device and instantiation; engines on several host threads share the flags
        static std::once_flag configured[64];
        cudaError_t err = cudaSuccess;
        std::call_once(configured[e->device & 63], [&] {
            err = cudaFuncSetAttribute(ntt_strided_kernel<D, GATHER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        });
        B2_CUDA(err);
    }
    ntt_strided_kernel<D, GATHER><<<grid, K1_THREADS, smem, e->stream>>>(p);
    B2_CHECK_LAUNCH();
    e->launches[0]++;
}
template <bool GATHER>
static void launch_strided(b2s_engine* e, unsigned D, const PassParams& p, dim3 grid) {
    switch (D) {
        case 1: launch_strided_t<1, GATHER>(e, p, grid); break;
        case 2: launch_strided_t<2, GATHER>(e, p, grid); break;
        case 3: launch_strided_t<3, GATHER>(e, p, grid); break;
        case 4: launch_strided_t<4, GATHER>(e, p, grid); break;
        case 5: launch_strided_t<5, GATHER>(e, p, grid); break;
        case 6: launch_strided_t<6, GATHER>(e, p, grid); break;
        case 7: launch_strided_t<7, GATHER>(e, p, grid); break;
        case 8: launch_strided_t<8, GATHER>(e, p, grid); break;
        case 9: launch_strided_t<9, GATHER>(e, p, grid); break;
        default: fail(B2S_ERR_INTERNAL, "ntt: unsupported strided digit");
    }
}
template <int D>
static void launch_contig_t(b2s_engine* e, const PassParams& p, unsigned cosets) {
    using C = K2Cfg<D>;
    const uint64_t total_runs = (uint64_t)p.width << (p.log_len - D);
    const uint64_t blocks = (total_runs + C::RPC - 1) / C::RPC;
    B2_REQUIRE(blocks <= 0x7FFFFFFFull, B2S_ERR_SHAPE, "ntt: too many runs in one launch");
    constexpr size_t smem = (size_t)C::RPC * (C::PITCH + (2 << D)) * 4;   // runs + one twiddle set (w, w') per run slot
    if (smem > 48 * 1024) {
        static std::once_flag configured[64];
        cudaError_t err = cudaSuccess;
        std::call_once(configured[e->device & 63], [&] {
            err = cudaFuncSetAttribute(ntt_contig_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        });
        B2_CUDA(err);
    }
    ntt_contig_kernel<D><<<dim3((unsigned)blocks, 1, cosets), C::THREADS, smem, e->stream>>>(p);
    B2_CHECK_LAUNCH();
    e->launches[0]++;
}
static void launch_contig(b2s_engine* e, unsigned D, const PassParams& p, unsigned cosets) {
    switch (D) {
        case 1: launch_contig_t<1>(e, p, cosets); break;
        case 2: launch_contig_t<2>(e, p, cosets); break;
        case 3: launch_contig_t<3>(e, p, cosets); break;
        case 4: launch_contig_t<4>(e, p, cosets); break;
        case 5: launch_contig_t<5>(e, p, cosets); break;
        case 6: launch_contig_t<6>(e, p, cosets); break;
        case 7: launch_contig_t<7>(e, p, cosets); break;
        case 8: launch_contig_t<8>(e, p, cosets); break;
        case 9: launch_contig_t<9>(e, p, cosets); break;
        case 10: launch_contig_t<10>(e, p, cosets); break;
        case 11: launch_contig_t<11>(e, p, cosets); break;
        case 12: launch_contig_t<12>(e, p, cosets); break;
        case 13: launch_contig_t<13>(e, p, cosets); break;
        default: fail(B2S_ERR_INTERNAL, "ntt: unsupported contiguous digit");
    }
}

// split the n depths into strided digits (<= 9 bits, top of the index) and one contiguous digit (<= 13 bits)
struct Plan {
    std::vector<unsigned> strided;
    unsigned contig;
};
static Plan make_plan(unsigned n) {
    Plan pl;
    // B2S_NTT_PLAN="n:d1,d2,...,dk" (tuning aid): strided digits d1.. and the contiguous digit dk for log length n
    if (const char* env = getenv("B2S_NTT_PLAN")) {
        unsigned en = 0;
        int used = 0;
        if (sscanf(env, "%u:%n", &en, &used) == 1 && en == n) {
            std::vector<unsigned> d;
            const char* q = env + used;
            while (*q) {
                char* end;
                unsigned v = (unsigned)strtoul(q, &end, 10);
                if (end == q) break;
                d.push_back(v);
                q = *end == ',' ? end + 1 : end;
            }
            unsigned sum = 0;
            bool ok = !d.empty() && d.back() >= 1 && d.back() <= 13;
            for (size_t i = 0; i < d.size(); i++) { sum += d[i]; if (i + 1 < d.size() && (d[i] < 1 || d[i] > K1_MAX_D)) ok = false; }
            // a strided pass needs at least 5 low bits below it (32 adjacent positions per tile row)
            if (ok && sum == n && d.back() >= 5) {
                pl.strided.assign(d.begin(), d.end() - 1);
                pl.contig = d.back();
                return pl;
            }
            fail(B2S_ERR_BAD_ARG, "B2S_NTT_PLAN: digits must sum to the log length (strided 1..9, contiguous 5..13)");
        }
    }
    // Every round of a pass is one trip through shared memory around a radix-2^R register block, R <= 4: the cost of a
    // plan is its number of rounds, sum of ceil(d / 4) (measured, tools/gpu_plan_sweep.sh: n = 20 as 8 + 12 = 5 rounds
    // runs in 5.98 ms, 9 + 11 = 6 rounds in 7.46 ms; n = 24 as 8 + 8 + 8 beats 6 + 6 + 12). Fewest passes first, then
    // fewest rounds, then the most even digits.
    auto rounds = [](unsigned d) { return (d + 3) / 4; };
    if (n <= 13) {
        pl.contig = n;
        return pl;
    }
    unsigned best_cost = ~0u, best_spread = ~0u;
    if (n <= K1_MAX_D + 13) {
        for (unsigned d1 = 4; d1 <= K1_MAX_D; d1++) {
            if (n < d1 + 5 || n - d1 > 13) continue;
            const unsigned d2 = n - d1, cost = rounds(d1) + rounds(d2), spread = d1 > d2 ? d1 - d2 : d2 - d1;
            if (cost < best_cost || (cost == best_cost && spread < best_spread)) {
                best_cost = cost; best_spread = spread;
                pl.strided = {d1};
                pl.contig = d2;
            }
        }
        return pl;
    }
    for (unsigned d1 = 4; d1 <= K1_MAX_D; d1++)
        for (unsigned d2 = 4; d2 <= d1; d2++) {
            if (n < d1 + d2 + 5 || n - d1 - d2 > 13) continue;
            const unsigned d3 = n - d1 - d2, cost = rounds(d1) + rounds(d2) + rounds(d3);
            const unsigned hi = std::max(d1, d3), lo = std::min(d2, d3), spread = hi - lo;
            if (cost < best_cost || (cost == best_cost && spread < best_spread)) {
                best_cost = cost; best_spread = spread;
                pl.strided = {d1, d2};
                pl.contig = d3;
            }
        }
    return pl;
}

// LDE of `width` columns: in = evaluations over shift_in*H_N, natural order (column stride in_col_stride);
// out = evaluations over GENERATOR*H_{N*2^b}, bit-reversed order (column stride out_col_stride >= N*2^b).
void lde_columns(b2s_engine* e, const uint32_t* in, uint64_t in_col_stride, unsigned logn, uint64_t width,
                 unsigned log_blowup, Fp shift_in, uint32_t* out, uint64_t out_col_stride, const PeerScatter* ps) {
    B2_REQUIRE(logn + log_blowup <= TWO_ADICITY, B2S_ERR_SHAPE, "lde: log height + log blowup exceeds two-adicity 27");
    if (width == 0) return;
    const uint64_t N = (uint64_t)1 << logn;
    const uint32_t B = 1u << log_blowup;
    B2_REQUIRE(!ps || logn >= 2, B2S_ERR_SHAPE, "lde: a row-sharded output needs at least 4 rows");
    if (logn == 0) {
        for (uint64_t c0 = 0; c0 < width; c0 += 1u << 20) {
            uint32_t g = (uint32_t)std::min<uint64_t>(width - c0, 1u << 20);
            replicate_kernel<<<(g * B + 255) / 256, 256, 0, e->stream>>>(in + c0 * in_col_stride, in_col_stride,
                                                                         out + c0 * out_col_stride, out_col_stride, g, B);
            B2_CHECK_LAUNCH();
            e->launches[0]++;
        }
        return;
    }
    TwiddleSet* tw = get_twiddles(e, logn);
    ScaleSet* sc = get_scales(e, logn, log_blowup, shift_in);
    const Plan pl = make_plan(logn);

    // column group: bounded coefficient workspace (2 GiB). Wider groups share each sub-problem's twiddles between more
    // columns: 2^24 x 256 LDE 88.8 ms with 2^27-element groups, 83.7 ms with 2^29 (profiles/r2g_lde_shoup.md)
    uint64_t group_elems = (uint64_t)1 << 29;
    if (const char* env = getenv("B2S_LDE_GROUP_ELEMS")) group_elems = strtoull(env, nullptr, 10);
    uint64_t G = group_elems / N;
    if (G < 1) G = 1;
    if (G > width) G = width;
    if (G > 65535) G = 65535;
    uint32_t* coef = (uint32_t*)e->dmalloc(G * N * 4);
    // row-sharded output: the passes before the last work in a group-sized scratch, the last one stores to the peers
    uint32_t* scratch = ps ? (uint32_t*)e->dmalloc(G * N * B * 4) : nullptr;
    if (ps) out_col_stride = N * B;

    for (uint64_t c0 = 0; c0 < width; c0 += G) {
        const uint32_t g = (uint32_t)std::min<uint64_t>(G, width - c0);
        const uint32_t* gin = in + c0 * in_col_stride;
        uint32_t* gout = ps ? scratch : out + c0 * out_col_stride;
        auto to_peers = [&](PassParams& q) {
            if (!ps) return;
            q.peer_n = ps->n_peers;
            q.peer_log_rows = ps->log_rows;
            q.peer_col0 = ps->col0 + c0;
            for (uint32_t i = 0; i < ps->n_peers; i++) q.peer_out[i] = ps->peer[i];
        };
        // ---------------- inverse: evaluations (natural) -> coefficients at bit-reversed slots, unscaled
        PassParams p{};
        p.width = g;
        p.log_len = logn;
        p.z = tw->zi;
        p.in = gin; p.in_col_stride = in_col_stride;
        p.out = coef; p.out_col_stride = N;
        unsigned t0 = 0;
        for (unsigned d : pl.strided) {
            p.t0 = t0;
            launch_strided<false>(e, d, p, dim3((unsigned)(N >> (d + 5)), g, 1));
            p.in = coef; p.in_col_stride = N;
            t0 += d;
        }
        p.t0 = t0;
        launch_contig(e, pl.contig, p, 1);
        // ---------------- forward on every coset: coefficients -> bit-reversed evaluations, block bitrev_b(coset)
        p = PassParams{};
        p.width = g;
        p.log_len = logn;
        p.z = sc->zf;
        p.log_blowup = log_blowup;
        p.coset = 1;
        p.in = coef; p.in_col_stride = N;
        p.out = gout; p.out_col_stride = out_col_stride;
        p.scale = sc->ninv;
        if (pl.strided.empty()) {
            p.t0 = 0;
            p.reverse_load = 1;
            to_peers(p);
            launch_contig(e, pl.contig, p, B);
        } else {
            p.t0 = 0;
            launch_strided<true>(e, pl.strided[0], p, dim3((unsigned)(N >> (pl.strided[0] + 5)), g, B));
            p.in = gout; p.in_col_stride = out_col_stride; p.in_block = 1; p.scale = 0;
            t0 = pl.strided[0];
            for (size_t i = 1; i < pl.strided.size(); i++) {
                p.t0 = t0;
                launch_strided<false>(e, pl.strided[i], p, dim3((unsigned)(N >> (pl.strided[i] + 5)), g, B));
                t0 += pl.strided[i];
            }
            p.t0 = t0;
            to_peers(p);
            launch_contig(e, pl.contig, p, B);
        }
    }
    if (scratch) e->dfree(scratch, G * N * B * 4);
    e->dfree(coef, G * N * 4);
}

}  // namespace b2
