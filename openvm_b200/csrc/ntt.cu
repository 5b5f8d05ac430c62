// BabyBear coset low-degree extension: inverse NTT -> coset scaling -> forward NTT with bit-reversed rows.
//
// Replaces (reference, external): the per-matrix part of Plonky3 `TwoAdicFriPcs::commit`
// (p3-fri / p3-dft `Radix2DitParallel::coset_lde_batch`), reached from
// crates/vm/src/arch/vm.rs:735 and crates/vm/src/system/program/trace.rs:58-63. Output convention pinned
// in-tree: leaf i of an LDE of height 2^h is the evaluation at GENERATOR * w_h^{bitrev_h(i)}
// (extensions/native/recursion/src/fri/two_adic_pcs.rs:178-186; domain: fri/domain.rs:142-154).
//
// Design (B200): a column of N = 2^n evaluations is transformed in at most three HBM passes per transform.
// n is split into digits of <= 9 bits; each pass stages a D x T tile (8192 words) in shared memory with
// >= 64-byte coalesced runs, runs the D-point transform as radix-16 register butterflies with shared-memory
// exchanges, applies the inter-digit twiddle from a two-level table and stores coalesced.
//   inverse:  A (strided)  ->  B (strided, in place)  ->  C (transposing; emits natural-order coefficients)
//   forward:  F1 (strided; fused coset scaling s_j^i / N, one launch for all 2^b cosets)
//             -> F2 (strided, in place) -> F3 (contiguous, in place); decimation-in-frequency, so the
//             result lands in bit-reversed order with no separate permutation pass. Coset j of the
//             2^b*N-point domain is exactly output block bitrev_b(j) (rows [blk*N, (blk+1)*N)).
#include "common.cuh"

namespace b2 {

constexpr int NTT_THREADS = 256;
constexpr int LOG_TILE = 13;  // words per tile
constexpr int MAX_DIGIT = 9;

// w_16^e and w_16^-e, e = 0..7, Montgomery form
static __constant__ uint32_t W16_FWD[8];
static __constant__ uint32_t W16_INV[8];

struct TwoLevel {       // x^e = lo[e & ((1<<h1)-1)] * hi[e >> h1]
    const uint32_t* lo;
    const uint32_t* hi;
    uint32_t h1;
};
B2_D Fp two_level(const TwoLevel& t, uint32_t e) {
    return Fp::raw(__ldg(t.lo + (e & ((1u << t.h1) - 1)))) * Fp::raw(__ldg(t.hi + (e >> t.h1)));
}

enum { MODE_STRIDED = 0, MODE_CONTIG = 1, MODE_TRANSPOSE = 2 };

struct PassParams {
    const uint32_t* in;
    uint32_t* out;
    uint64_t in_col_stride, out_col_stride;
    uint32_t width;       // columns
    uint32_t mode;
    uint32_t logT;
    uint32_t logL;        // strided: log2 of the run length below the digit; transpose: log2 E1 (middle digit)
    uint32_t logE0;       // transpose: log2 of the leading digit (becomes the lowest coefficient digit)
    uint32_t log_len;     // log2 of the per-column length covered by this launch
    uint32_t out_brev;    // 1: frequency k stored at slot bitrev(k) (DIF order); 0: at slot k
    uint32_t inverse;     // selects W16_INV
    TwoLevel tw;          // root table of order 2^tw_logn (forward or inverse)
    uint32_t tw_logn;
    uint32_t tw_shift;    // strided: exponent = (lo*k) << tw_shift
    uint32_t has_scale;   // multiply loaded element i (index in column, mod 2^sc_logn) by scale^i
    TwoLevel sc;          // tables for coset blockIdx.z start at sc.lo + z*sc_lo_stride etc.
    uint32_t sc_lo_stride, sc_hi_stride;
    uint32_t log_blowup;  // out block of coset z is bitrev_b(z), block length 2^log_len
};

template <int R>
B2_D void reg_ntt(Fp (&v)[1 << R], const uint32_t* __restrict__ w16) {
    // decimation in frequency, natural in, bit-reversed out; twiddles w_{2^R}^x = w_16^{x * 16 / 2^R}
#pragma unroll
    for (int s = 0; s < R; s++) {
        const int half = 1 << (R - 1 - s);
#pragma unroll
        for (int blk = 0; blk < (1 << R); blk += 2 * half) {
#pragma unroll
            for (int j = 0; j < half; j++) {
                Fp u = v[blk + j], w = v[blk + j + half];
                v[blk + j] = u + w;
                const int e = (j << s) * (16 >> R);  // exponent of w_16
                if (e == 0) {
                    v[blk + j + half] = u - w;
                } else {
                    // (u - w + p) in (0, 2p): valid lazy input to the Montgomery product
                    uint32_t d = u.v - w.v + P;
                    v[blk + j + half] = Fp::raw(monty_reduce((uint64_t)d * w16[e]));
                }
            }
        }
    }
}

template <int LOG_D, int REM>
struct Rounds {
    static B2_D void run(uint32_t* x, const uint32_t* wd, uint32_t logT, uint32_t pitch, const uint32_t* w16) {
        constexpr int R = REM >= 4 ? 4 : REM;
        constexpr int SH = REM - R;
        const uint32_t T = 1u << logT;
        const uint32_t items = (1u << (LOG_D - R)) << logT;
        for (uint32_t id = threadIdx.x; id < items; id += NTT_THREADS) {
            uint32_t t = id & (T - 1);
            uint32_t g = id >> logT;
            uint32_t r = g & ((1u << SH) - 1);
            uint32_t d0 = ((g >> SH) << (SH + R)) + r;
            uint32_t* px = x + d0 * pitch + t;
            Fp v[1 << R];
#pragma unroll
            for (int a = 0; a < (1 << R); a++) v[a] = Fp::raw(px[(a << SH) * pitch]);
            reg_ntt<R>(v, w16);
            if (SH > 0) {
                // slot a holds sub-frequency ka = bitrev_R(a); twiddle w_{2^(SH+R)}^{r*ka}
#pragma unroll
                for (int a = 1; a < (1 << R); a++) {
                    const uint32_t ka = __brev((uint32_t)a) >> (32 - R);
                    v[a] = v[a] * Fp::raw(wd[(r * ka) << (LOG_D - SH - R)]);
                }
            }
#pragma unroll
            for (int a = 0; a < (1 << R); a++) px[(a << SH) * pitch] = v[a].v;
        }
        __syncthreads();
        Rounds<LOG_D, SH>::run(x, wd, logT, pitch, w16);
    }
};
template <int LOG_D>
struct Rounds<LOG_D, 0> {
    static B2_D void run(uint32_t*, const uint32_t*, uint32_t, uint32_t, const uint32_t*) {}
};

template <int LOG_D>
__global__ void __launch_bounds__(NTT_THREADS) ntt_pass_kernel(PassParams p) {
    extern __shared__ uint32_t sm[];
    constexpr uint32_t D = 1u << LOG_D;
    const uint32_t logT = p.logT, T = 1u << logT;
    const uint32_t pitch = T == 1 ? 1 : T + 1;
    uint32_t* x = sm;
    uint32_t* wd = sm + D * pitch;
    const uint32_t* w16 = p.inverse ? W16_INV : W16_FWD;
    const uint32_t tid = threadIdx.x;

    for (uint32_t e = tid; e < D; e += NTT_THREADS) wd[e] = two_level(p.tw, e << (p.tw_logn - LOG_D)).v;

    // ---- tile coordinates ----
    const uint64_t tile = blockIdx.x;
    const uint32_t z = blockIdx.z;
    uint64_t in_base = 0, out_base = 0;       // word offsets of element (d=0, t=0)
    uint64_t in_sd = 1, in_st = 1, out_sd = 1, out_st = 1;
    uint32_t lo0 = 0;                         // strided: index within the run of the tile's t = 0
    uint64_t sub0 = 0, total_subs = 0;        // contiguous: first sub-array of the tile, count over all columns
    const uint32_t logS = p.log_len - LOG_D;  // sub-arrays per column (contiguous mode)
    if (p.mode == MODE_STRIDED) {
        const uint32_t col = blockIdx.y;
        const uint32_t chunks = 1u << (p.logL - logT);          // chunks of T per run
        const uint64_t hi = tile >> (p.logL - logT);
        lo0 = (uint32_t)(tile & (chunks - 1)) << logT;
        uint64_t pos = (hi << (LOG_D + p.logL)) + lo0;
        in_base = (uint64_t)col * p.in_col_stride + pos;
        out_base = (uint64_t)col * p.out_col_stride + pos;
        in_sd = out_sd = (uint64_t)1 << p.logL;
    } else if (p.mode == MODE_TRANSPOSE) {
        // in (i0, i1, d) -> out (k, i1, i0); tile = (chunk of T values of i0, i1)
        const uint32_t col = blockIdx.y;
        const uint32_t logE1 = p.logL;
        const uint64_t i1 = tile & ((1u << logE1) - 1);
        const uint64_t i0 = (tile >> logE1) << logT;
        in_base = (uint64_t)col * p.in_col_stride + (i0 << (logE1 + LOG_D)) + (i1 << LOG_D);
        in_st = (uint64_t)1 << (logE1 + LOG_D);
        out_base = (uint64_t)col * p.out_col_stride + (i1 << p.logE0) + i0;
        out_sd = (uint64_t)1 << (p.logE0 + logE1);
    } else {
        sub0 = tile << logT;
        total_subs = (uint64_t)p.width << logS;
    }
    // coset z of the extended domain is output block bitrev_b(z)
    const uint64_t blk_off = (p.has_scale && p.log_blowup) ? ((uint64_t)(__brev(z) >> (32 - p.log_blowup)) << p.log_len) : 0;
    out_base += blk_off;
    TwoLevel sc = p.sc;
    if (p.has_scale) { sc.lo += (uint64_t)z * p.sc_lo_stride; sc.hi += (uint64_t)z * p.sc_hi_stride; }
    const uint32_t len_mask = (p.log_len >= 32) ? 0xFFFFFFFFu : ((1u << p.log_len) - 1);

    // ---- load ----
    const uint32_t E = D << logT;
    if (p.mode == MODE_STRIDED) {
        for (uint32_t e = tid; e < E; e += NTT_THREADS) {
            uint32_t t = e & (T - 1), d = e >> logT;
            uint64_t off = (uint64_t)d * in_sd + t;
            Fp v = Fp::raw(p.in[in_base + off]);
            if (p.has_scale) {
                uint32_t i = (uint32_t)(((tile >> (p.logL - logT)) << (LOG_D + p.logL)) + lo0 + off) & len_mask;
                v = v * two_level(sc, i);
            }
            x[d * pitch + t] = v.v;
        }
    } else if (p.mode == MODE_TRANSPOSE) {
        for (uint32_t e = tid; e < E; e += NTT_THREADS) {
            uint32_t d = e & (D - 1), t = e >> LOG_D;
            x[d * pitch + t] = p.in[in_base + (uint64_t)t * in_st + d];
        }
    } else {
        for (uint32_t e = tid; e < E; e += NTT_THREADS) {
            uint32_t d = e & (D - 1), t = e >> LOG_D;
            uint64_t g = sub0 + t;
            uint32_t v = 0;
            if (g < total_subs) {
                uint64_t col = g >> logS, sub = g & (((uint64_t)1 << logS) - 1);
                uint64_t i = (sub << LOG_D) + d;
                Fp f = Fp::raw(p.in[col * p.in_col_stride + i]);
                if (p.has_scale) f = f * two_level(sc, (uint32_t)i & len_mask);
                v = f.v;
            }
            x[d * pitch + t] = v;
        }
    }
    __syncthreads();

    Rounds<LOG_D, LOG_D>::run(x, wd, logT, pitch, w16);

    // ---- store: smem position q holds frequency bitrev(q) ----
    if (p.mode == MODE_CONTIG) {
        for (uint32_t e = tid; e < E; e += NTT_THREADS) {
            uint32_t slot = e & (D - 1), t = e >> LOG_D;
            uint64_t g = sub0 + t;
            if (g >= total_subs) continue;
            uint32_t q = p.out_brev ? slot : (LOG_D ? (__brev(slot) >> (32 - LOG_D)) : 0);
            uint64_t col = g >> logS, sub = g & (((uint64_t)1 << logS) - 1);
            p.out[col * p.out_col_stride + blk_off + (sub << LOG_D) + slot] = x[q * pitch + t];
        }
    } else {
        for (uint32_t e = tid; e < E; e += NTT_THREADS) {
            uint32_t t = e & (T - 1), slot = e >> logT;
            uint32_t b = LOG_D ? (__brev(slot) >> (32 - LOG_D)) : 0;
            uint32_t q = p.out_brev ? slot : b;
            uint32_t k = p.out_brev ? b : slot;
            Fp v = Fp::raw(x[q * pitch + t]);
            if (p.mode == MODE_STRIDED && k) v = v * two_level(p.tw, ((lo0 + t) * k) << p.tw_shift);
            p.out[out_base + (uint64_t)slot * out_sd + (uint64_t)t * out_st] = v.v;
        }
    }
}

struct TwiddleSet {
    unsigned logn;
    uint32_t h1;
    uint32_t *fw_lo, *fw_hi, *inv_lo, *inv_hi;
    size_t bytes;
};

// scale tables for (logn, log_blowup, shift_in): for coset j: (GENERATOR/shift_in * w_M^j)^i / N
struct ScaleSet {
    uint32_t h1;
    uint32_t* lo;  // [2^b][2^h1]
    uint32_t* hi;  // [2^b][2^(logn-h1)]
    size_t bytes;
};

static void build_two_level(Fp base, unsigned logn, uint32_t h1, Fp hi_factor, std::vector<uint32_t>& lo,
                            std::vector<uint32_t>& hi) {
    uint32_t nlo = 1u << h1, nhi = 1u << (logn - h1);
    lo.resize(nlo);
    hi.resize(nhi);
    Fp a = Fp::one();
    for (uint32_t i = 0; i < nlo; i++) { lo[i] = a.v; a *= base; }
    Fp step = a;  // base^(2^h1)
    Fp b = hi_factor;
    for (uint32_t i = 0; i < nhi; i++) { hi[i] = b.v; b *= step; }
}

static uint32_t* upload(b2s_engine* e, const std::vector<uint32_t>& v) {
    uint32_t* d = (uint32_t*)e->dmalloc(v.size() * 4);
    B2_CUDA(cudaMemcpyAsync(d, v.data(), v.size() * 4, cudaMemcpyHostToDevice, e->stream));
    B2_CUDA(cudaStreamSynchronize(e->stream));
    return d;
}

static TwiddleSet* get_twiddles(b2s_engine* e, unsigned logn) {
    auto it = e->twiddles.find(logn);
    if (it != e->twiddles.end()) return it->second;
    if (e->twiddles.empty()) {
        uint32_t f[8], g[8];
        Fp w = two_adic_generator(4), wi = fp_inv(w), a = Fp::one(), b = Fp::one();
        for (int i = 0; i < 8; i++) { f[i] = a.v; g[i] = b.v; a *= w; b *= wi; }
        B2_CUDA(cudaMemcpyToSymbol(W16_FWD, f, sizeof(f)));
        B2_CUDA(cudaMemcpyToSymbol(W16_INV, g, sizeof(g)));
    }
    TwiddleSet* t = new TwiddleSet();
    t->logn = logn;
    t->h1 = (logn + 1) / 2;
    std::vector<uint32_t> lo, hi;
    Fp w = two_adic_generator(logn);
    build_two_level(w, logn, t->h1, Fp::one(), lo, hi);
    t->fw_lo = upload(e, lo);
    t->fw_hi = upload(e, hi);
    build_two_level(fp_inv(w), logn, t->h1, Fp::one(), lo, hi);
    t->inv_lo = upload(e, lo);
    t->inv_hi = upload(e, hi);
    t->bytes = (lo.size() + hi.size()) * 8;
    e->twiddles[logn] = t;
    return t;
}

void free_twiddles(b2s_engine* e) {
    for (auto& kv : e->twiddles) {
        TwiddleSet* t = kv.second;
        size_t nlo = (size_t)4 << t->h1, nhi = (size_t)4 << (t->logn - t->h1);
        e->dfree(t->fw_lo, nlo); e->dfree(t->fw_hi, nhi); e->dfree(t->inv_lo, nlo); e->dfree(t->inv_hi, nhi);
        delete t;
    }
    e->twiddles.clear();
}

template <int LOG_D>
static void launch_pass_t(b2s_engine* e, const PassParams& p, dim3 grid) {
    const uint32_t D = 1u << LOG_D, T = 1u << p.logT;
    const uint32_t pitch = T == 1 ? 1 : T + 1;
    size_t smem = (size_t)(D * pitch + D) * 4;
    ntt_pass_kernel<LOG_D><<<grid, NTT_THREADS, smem, e->stream>>>(p);
    B2_CHECK_LAUNCH();
    e->launches[0]++;
}

static void launch_pass(b2s_engine* e, unsigned logD, const PassParams& p, dim3 grid) {
    switch (logD) {
        case 0: launch_pass_t<0>(e, p, grid); break;
        case 1: launch_pass_t<1>(e, p, grid); break;
        case 2: launch_pass_t<2>(e, p, grid); break;
        case 3: launch_pass_t<3>(e, p, grid); break;
        case 4: launch_pass_t<4>(e, p, grid); break;
        case 5: launch_pass_t<5>(e, p, grid); break;
        case 6: launch_pass_t<6>(e, p, grid); break;
        case 7: launch_pass_t<7>(e, p, grid); break;
        case 8: launch_pass_t<8>(e, p, grid); break;
        case 9: launch_pass_t<9>(e, p, grid); break;
        default: fail(B2S_ERR_INTERNAL, "ntt: unsupported digit size");
    }
}

// split n into k = ceil(n/9) digits, as even as possible, largest first
static std::vector<unsigned> split_digits(unsigned n) {
    if (n == 0) return {0};
    unsigned k = (n + MAX_DIGIT - 1) / MAX_DIGIT;
    std::vector<unsigned> d(k, n / k);
    for (unsigned i = 0; i < n % k; i++) d[i]++;
    return d;
}


static ScaleSet* get_scales(b2s_engine* e, unsigned logn, unsigned log_blowup, Fp shift_in) {
    uint64_t key = ((uint64_t)logn << 40) | ((uint64_t)log_blowup << 32) | shift_in.v;
    auto it = e->scales.find(key);
    if (it != e->scales.end()) return it->second;
    ScaleSet* s = new ScaleSet();
    s->h1 = (logn + 1) / 2;
    const uint32_t B = 1u << log_blowup, nlo = 1u << s->h1, nhi = 1u << (logn - s->h1);
    std::vector<uint32_t> lo_all((size_t)B * nlo), hi_all((size_t)B * nhi), lo, hi;
    Fp base = Fp::raw(M_GENERATOR) * fp_inv(shift_in);
    Fp wm = two_adic_generator(logn + log_blowup);
    Fp ninv = fp_inv(Fp::from_canonical((uint32_t)(((uint64_t)1 << logn) % P)));
    Fp sj = base;
    for (uint32_t j = 0; j < B; j++) {
        build_two_level(sj, logn, s->h1, ninv, lo, hi);
        memcpy(&lo_all[(size_t)j * nlo], lo.data(), (size_t)nlo * 4);
        memcpy(&hi_all[(size_t)j * nhi], hi.data(), (size_t)nhi * 4);
        sj *= wm;
    }
    s->lo = upload(e, lo_all);
    s->hi = upload(e, hi_all);
    s->bytes = (lo_all.size() + hi_all.size()) * 4;
    e->scales[key] = s;
    return s;
}

void free_scales(b2s_engine* e) {
    for (auto& kv : e->scales) {
        ScaleSet* s = kv.second;
        unsigned logn = (unsigned)(kv.first >> 40), lb = (unsigned)((kv.first >> 32) & 0xFF);
        e->dfree(s->lo, ((size_t)4 << s->h1) << lb);
        e->dfree(s->hi, ((size_t)4 << (logn - s->h1)) << lb);
        delete s;
    }
    e->scales.clear();
}

static uint32_t pick_logT(unsigned logD, unsigned log_run) {
    unsigned t = LOG_TILE - logD;
    return t < log_run ? t : log_run;
}

// LDE of `width` columns: in = evaluations over shift_in*H_N, natural order (column stride in_col_stride);
// out = evaluations over GENERATOR*H_{N*2^b}, bit-reversed order (column stride out_col_stride >= N*2^b).
void lde_columns(b2s_engine* e, const uint32_t* in, uint64_t in_col_stride, unsigned logn, uint64_t width,
                 unsigned log_blowup, Fp shift_in, uint32_t* out, uint64_t out_col_stride) {
    B2_REQUIRE(logn + log_blowup <= TWO_ADICITY, B2S_ERR_SHAPE, "lde: log height + log blowup exceeds two-adicity 27");
    if (width == 0) return;
    const uint64_t N = (uint64_t)1 << logn;
    const uint32_t B = 1u << log_blowup;
    TwiddleSet* tw = get_twiddles(e, logn);
    ScaleSet* sc = get_scales(e, logn, log_blowup, shift_in);
    std::vector<unsigned> dg = split_digits(logn);
    const size_t k = dg.size();
    TwoLevel tw_f{tw->fw_lo, tw->fw_hi, tw->h1}, tw_i{tw->inv_lo, tw->inv_hi, tw->h1};

    // column group size: bounded workspace (two N-word buffers per column in flight)
    uint64_t group_elems = (uint64_t)1 << 25;
    if (const char* env = getenv("B2S_LDE_GROUP_ELEMS")) group_elems = strtoull(env, nullptr, 10);
    uint64_t G = group_elems / N;
    if (G < 1) G = 1;
    if (G > width) G = width;
    if (G > 65535) G = 65535;
    uint32_t* ws1 = k > 1 ? (uint32_t*)e->dmalloc(G * N * 4) : nullptr;
    uint32_t* coef = (uint32_t*)e->dmalloc(G * N * 4);

    for (uint64_t c0 = 0; c0 < width; c0 += G) {
        const uint32_t g = (uint32_t)std::min<uint64_t>(G, width - c0);
        const uint32_t* gin = in + c0 * in_col_stride;
        uint32_t* gout = out + c0 * out_col_stride;
        PassParams p{};
        p.width = g;
        p.tw_logn = logn;
        // ---------------- inverse: evaluations -> natural-order coefficients (unscaled; 1/N is in the scale table)
        p.inverse = 1;
        p.tw = tw_i;
        p.has_scale = 0;
        p.out_brev = 0;
        if (k == 1) {
            p.mode = MODE_CONTIG; p.in = gin; p.out = coef; p.in_col_stride = in_col_stride; p.out_col_stride = N;
            p.log_len = logn; p.logT = LOG_TILE - dg[0];
            uint64_t subs = (uint64_t)g;  // one sub-array per column
            launch_pass(e, dg[0], p, dim3((unsigned)((subs + (1u << p.logT) - 1) >> p.logT), 1, 1));
        } else {
            // pass A
            p.mode = MODE_STRIDED; p.in = gin; p.out = ws1; p.in_col_stride = in_col_stride; p.out_col_stride = N;
            p.log_len = logn; p.logL = logn - dg[0]; p.logT = pick_logT(dg[0], p.logL); p.tw_shift = 0;
            launch_pass(e, dg[0], p, dim3((unsigned)(N >> (dg[0] + p.logT)), g, 1));
            if (k == 3) {
                p.in = ws1; p.in_col_stride = N;
                p.logL = dg[2]; p.logT = pick_logT(dg[1], p.logL); p.tw_shift = dg[0];
                launch_pass(e, dg[1], p, dim3((unsigned)(N >> (dg[1] + p.logT)), g, 1));
            }
            // pass C
            p.mode = MODE_TRANSPOSE; p.in = ws1; p.out = coef; p.in_col_stride = N; p.out_col_stride = N;
            p.logE0 = dg[0]; p.logL = (k == 3) ? dg[1] : 0;
            p.logT = pick_logT(dg[k - 1], dg[0]);
            launch_pass(e, dg[k - 1], p, dim3((unsigned)(N >> (dg[k - 1] + p.logT)), g, 1));
        }
        // ---------------- forward on each coset: digits in reverse order
        p = PassParams{};
        p.width = g;
        p.tw_logn = logn;
        p.inverse = 0;
        p.tw = tw_f;
        p.out_brev = 1;
        p.has_scale = 1;
        p.sc = TwoLevel{sc->lo, sc->hi, sc->h1};
        p.sc_lo_stride = 1u << sc->h1;
        p.sc_hi_stride = 1u << (logn - sc->h1);
        p.log_blowup = log_blowup;
        p.in = coef; p.in_col_stride = N; p.out = gout; p.out_col_stride = out_col_stride;
        p.log_len = logn;
        if (k == 1) {
            p.mode = MODE_CONTIG; p.logT = LOG_TILE - dg[0];
            launch_pass(e, dg[0], p, dim3((unsigned)(((uint64_t)g + (1u << p.logT) - 1) >> p.logT), 1, B));
        } else {
            const unsigned f0 = dg[k - 1];
            p.mode = MODE_STRIDED; p.logL = logn - f0; p.logT = pick_logT(f0, p.logL); p.tw_shift = 0;
            launch_pass(e, f0, p, dim3((unsigned)(N >> (f0 + p.logT)), g, B));
            // remaining passes run in place over the whole extended column (2^b blocks of N)
            p.has_scale = 0; p.in = gout; p.in_col_stride = out_col_stride; p.log_len = logn + log_blowup;
            const uint64_t M = N << log_blowup;
            if (k == 3) {
                p.logL = dg[0]; p.logT = pick_logT(dg[1], p.logL); p.tw_shift = f0;
                launch_pass(e, dg[1], p, dim3((unsigned)(M >> (dg[1] + p.logT)), g, 1));
            }
            p.mode = MODE_CONTIG; p.logT = LOG_TILE - dg[0];
            uint64_t subs = (uint64_t)g << (logn + log_blowup - dg[0]);
            launch_pass(e, dg[0], p, dim3((unsigned)((subs + (1u << p.logT) - 1) >> p.logT), 1, 1));
        }
    }
    if (ws1) e->dfree(ws1, G * N * 4);
    e->dfree(coef, G * N * 4);
}

}  // namespace b2
