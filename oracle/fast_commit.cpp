// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header). Never linked into or called by the product.
//
// orc_fast_commit: the commitment of ONE trace matrix (coset LDE + Poseidon2 Merkle root; the same function as
// orc_commit on one matrix), computed the way the reference's CPU prover computes it -- Plonky3's rayon path with
// packed AVX2 field arithmetic (p3-baby-bear `PackedBabyBearAVX2`, p3-dft `Radix2DitParallel`, p3-merkle-tree with
// packed Poseidon2; absent dependency @ 539bbc84, reached from crates/vm/src/arch/vm.rs:735) -- instead of the plain
// oracle's scalar '%' arithmetic:
//   * Montgomery form, 8 lanes per AVX2 vector (even / odd 32x32->64 products);
//   * radix-2 decimation-in-frequency butterflies, per-stage twiddle tables, the low stages blocked for L2;
//     one column per thread;
//   * leaf sponge and 2-to-1 compression vectorised over 8 adjacent rows / nodes; rows split over the threads.
// It exists for two jobs (bench.py, tests): the CPU baseline at the benchmark's own size, and the check of the
// 2^24 x 256 root, both of which the plain oracle is too slow for. tests/test_oracle.py pins it to the plain oracle
// (orc_commit, independent canonical arithmetic) on small shapes; the golden vectors pin the plain oracle.
//
// Domain conventions restated (extensions/native/recursion/src/fri/two_adic_pcs.rs:178-186): row i of the LDE of
// height M = N 2^b holds f(GENERATOR * w_M^{bitrev(i)}); hence coset j (points GENERATOR w_M^j w_N^k) is the block of N
// rows starting at bitrev_b(j) N, in bit-reversed order of k.
#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

#include "poseidon2.hpp"   // constants (canonical) from the plain oracle's generator; arithmetic below is independent

namespace fastc {

using orc::P;
static uint32_t NEG_INV;   // -P^-1 mod 2^32
static uint32_t R2;        // 2^64 mod P

static inline uint32_t mmul(uint32_t a, uint32_t b) {
    const uint64_t t = (uint64_t)a * b;
    const uint32_t m = (uint32_t)t * NEG_INV;
    const uint32_t r = (uint32_t)((t + (uint64_t)m * P) >> 32);
    return r >= P ? r - P : r;
}
static inline uint32_t to_m(uint32_t a) { return mmul(a, R2); }
static inline uint32_t from_m(uint32_t a) { return mmul(a, 1); }
static inline uint32_t madd(uint32_t a, uint32_t b) { uint32_t s = a + b; return s >= P ? s - P : s; }
static inline uint32_t msub(uint32_t a, uint32_t b) { return a >= b ? a - b : a + P - b; }
static uint32_t mpow(uint32_t b, uint64_t e) {
    uint32_t r = to_m(1);
    while (e) { if (e & 1) r = mmul(r, b); b = mmul(b, b); e >>= 1; }
    return r;
}

typedef __m256i V;
static inline V vset(uint32_t x) { return _mm256_set1_epi32((int)x); }
static inline V vadd(V a, V b) { V s = _mm256_add_epi32(a, b); return _mm256_min_epu32(s, _mm256_sub_epi32(s, vset(P))); }
static inline V vsub(V a, V b) { V d = _mm256_sub_epi32(a, b); return _mm256_min_epu32(d, _mm256_add_epi32(d, vset(P))); }
static inline V vmul(V a, V b) {
    const V pv = vset(P), ni = vset(NEG_INV);
    const V te = _mm256_mul_epu32(a, b), to = _mm256_mul_epu32(_mm256_srli_epi64(a, 32), _mm256_srli_epi64(b, 32));
    const V me = _mm256_mul_epu32(te, ni), mo = _mm256_mul_epu32(to, ni);
    const V ue = _mm256_add_epi64(te, _mm256_mul_epu32(me, pv)), uo = _mm256_add_epi64(to, _mm256_mul_epu32(mo, pv));
    const V r = _mm256_blend_epi32(_mm256_srli_epi64(ue, 32), uo, 0xAA);
    return _mm256_min_epu32(r, _mm256_sub_epi32(r, pv));
}
static inline V vload(const uint32_t* p) { return _mm256_loadu_si256((const __m256i*)p); }
static inline void vstore(uint32_t* p, V v) { _mm256_storeu_si256((__m256i*)p, v); }

static void init_constants() {
    static bool done = false;
    if (done) return;
    uint32_t inv = 1;                                   // Newton: inv = P^-1 mod 2^32
    for (int i = 0; i < 5; i++) inv *= 2 - P * inv;
    NEG_INV = 0u - inv;
    R2 = (uint32_t)((((uint64_t)1 << 32) % P) * (((uint64_t)1 << 32) % P) % P);
    done = true;
}

// ---- parallel helpers -------------------------------------------------------------------------------------------------
static void par_range(size_t n, int threads, const std::function<void(size_t, size_t)>& fn) {
    const size_t t = std::max<size_t>(1, std::min<size_t>((size_t)threads, n));
    if (t == 1) { fn(0, n); return; }
    std::vector<std::thread> th;
    const size_t chunk = (n + t - 1) / t;
    for (size_t i = 0; i < t; i++) {
        const size_t lo = i * chunk, hi = std::min(n, lo + chunk);
        if (lo < hi) th.emplace_back(fn, lo, hi);
    }
    for (auto& x : th) x.join();
}
static void par_items(size_t n, int threads, const std::function<void(size_t)>& fn) {   // dynamic: one item at a time
    std::atomic<size_t> next{0};
    const size_t t = std::max<size_t>(1, std::min<size_t>((size_t)threads, n));
    auto body = [&] { for (size_t i; (i = next.fetch_add(1)) < n;) fn(i); };
    if (t == 1) { body(); return; }
    std::vector<std::thread> th;
    for (size_t i = 0; i < t; i++) th.emplace_back(body);
    for (auto& x : th) x.join();
}

// ---- NTT: decimation in frequency, natural order in, bit-reversed order out --------------------------------------------
// table[half + j] = w_{2 half}^j (Montgomery) for half = 1, 2, 4, ... n/2
static std::vector<uint32_t> stage_table(unsigned logn, uint32_t root_m, int threads) {
    std::vector<uint32_t> t((size_t)1 << logn, 0);
    for (unsigned s = 1; s <= logn; s++) {
        const size_t half = (size_t)1 << (s - 1);
        uint32_t w = root_m;                            // root of order 2^logn -> order 2^s
        for (unsigned k = s; k < logn; k++) w = mmul(w, w);
        uint32_t* dst = t.data() + half;
        const size_t seg = 1 << 12;
        par_items((half + seg - 1) / seg, threads, [&](size_t q) {
            const size_t lo = q * seg, hi = std::min(half, lo + seg);
            uint32_t x = mpow(w, lo);
            for (size_t j = lo; j < hi; j++) { dst[j] = x; x = mmul(x, w); }
        });
    }
    return t;
}
static inline void dif_stage(uint32_t* a, size_t half, const uint32_t* tw) {   // one block of 2 * half words
    if (half >= 8) {
        for (size_t j = 0; j < half; j += 8) {
            const V u = vload(a + j), v = vload(a + j + half);
            vstore(a + j, vadd(u, v));
            vstore(a + j + half, vmul(vsub(u, v), vload(tw + j)));
        }
    } else {
        for (size_t j = 0; j < half; j++) {
            const uint32_t u = a[j], v = a[j + half];
            a[j] = madd(u, v);
            a[j + half] = mmul(msub(u, v), tw[j]);
        }
    }
}
static void dif_ntt(uint32_t* a, unsigned logn, const uint32_t* table) {
    const unsigned BLK = 15;                             // 2^15 words = 128 KB: the low stages stay in L2
    const size_t n = (size_t)1 << logn;
    unsigned s = logn;
    for (; s > BLK; s--) {
        const size_t half = (size_t)1 << (s - 1);
        for (size_t base = 0; base < n; base += 2 * half) dif_stage(a + base, half, table + half);
    }
    const size_t blk = (size_t)1 << s;
    for (size_t b0 = 0; b0 < n; b0 += blk)
        for (unsigned q = s; q >= 1; q--) {
            const size_t half = (size_t)1 << (q - 1);
            for (size_t base = b0; base < b0 + blk; base += 2 * half) dif_stage(a + base, half, table + half);
        }
}

static inline uint32_t brev(uint32_t x, unsigned bits) {
    static const uint8_t* R8 = [] {
        static uint8_t t[256];
        for (int i = 0; i < 256; i++) { int r = 0; for (int k = 0; k < 8; k++) r |= ((i >> k) & 1) << (7 - k); t[i] = (uint8_t)r; }
        return t;
    }();
    if (!bits) return 0;
    const uint32_t r = ((uint32_t)R8[x & 255] << 24) | ((uint32_t)R8[(x >> 8) & 255] << 16) | ((uint32_t)R8[(x >> 16) & 255] << 8) |
                       R8[x >> 24];
    return r >> (32 - bits);
}
// dst[i] = src[bitrev(i)] * scale[i]; tiles of 2^k x 2^k words keep both sides in cache for tall columns
static void bitrev_scale(uint32_t* dst, const uint32_t* src, const uint32_t* scale, unsigned logn) {
    const size_t n = (size_t)1 << logn;
    if (logn < 16) {
        for (size_t i = 0; i < n; i++) dst[i] = mmul(src[brev((uint32_t)i, logn)], scale[i]);
        return;
    }
    const unsigned k = 8, mid_bits = logn - 2 * k;
    for (uint32_t mid = 0; mid < (1u << mid_bits); mid++) {
        const uint32_t rmid = brev(mid, mid_bits);
        for (uint32_t hi = 0; hi < (1u << k); hi++) {
            const uint32_t* s = src + (((size_t)hi << (logn - k)) | ((size_t)mid << k));
            const uint32_t rhi = brev(hi, k);
            for (uint32_t lo = 0; lo < (1u << k); lo++) {
                const size_t d = ((size_t)brev(lo, k) << (logn - k)) | ((size_t)rmid << k) | rhi;
                dst[d] = mmul(s[lo], scale[d]);
            }
        }
    }
}

// ---- Poseidon2 on 8 states at once (structure: crates/circuits/poseidon2-air/src/permute.rs:104-146) ---------------------
struct P2Consts {
    uint32_t ext[8][16], partial[13], diag[16];          // Montgomery
};
static P2Consts g_p2;
static void init_p2() {
    const orc::Poseidon2Constants& c = orc::poseidon2_constants();
    for (int r = 0; r < 4; r++)
        for (int i = 0; i < 16; i++) { g_p2.ext[r][i] = to_m(c.ext_init[r][i].v); g_p2.ext[4 + r][i] = to_m(c.ext_term[r][i].v); }
    for (int r = 0; r < 13; r++) g_p2.partial[r] = to_m(c.partial[r].v);
    for (int i = 0; i < 16; i++) g_p2.diag[i] = to_m(c.diag[i].v);
}
static inline V vsbox(V x) { const V x2 = vmul(x, x), x3 = vmul(x2, x), x4 = vmul(x2, x2); return vmul(x3, x4); }
static inline void vext(V* s) {
    for (int b = 0; b < 16; b += 4) {
        const V a = s[b], bb = s[b + 1], c = s[b + 2], d = s[b + 3];
        const V ab = vadd(a, bb), cd = vadd(c, d), all = vadd(ab, cd);
        // [[2,3,1,1],[1,2,3,1],[1,1,2,3],[3,1,1,2]]: row k = (a + b + c + d) + x_k + 2 x_{k+1}
        s[b] = vadd(all, vadd(a, vadd(bb, bb)));
        s[b + 1] = vadd(all, vadd(bb, vadd(c, c)));
        s[b + 2] = vadd(all, vadd(c, vadd(d, d)));
        s[b + 3] = vadd(all, vadd(d, vadd(a, a)));
    }
    V sums[4];
    for (int k = 0; k < 4; k++) sums[k] = vadd(vadd(s[k], s[4 + k]), vadd(s[8 + k], s[12 + k]));
    for (int i = 0; i < 16; i++) s[i] = vadd(s[i], sums[i & 3]);
}
static inline void vperm(V* s) {
    vext(s);
    for (int r = 0; r < 4; r++) {
        for (int i = 0; i < 16; i++) s[i] = vsbox(vadd(s[i], vset(g_p2.ext[r][i])));
        vext(s);
    }
    for (int r = 0; r < 13; r++) {
        s[0] = vsbox(vadd(s[0], vset(g_p2.partial[r])));
        V sum = s[0];
        for (int i = 1; i < 16; i++) sum = vadd(sum, s[i]);
        for (int i = 0; i < 16; i++) s[i] = vadd(sum, vmul(s[i], vset(g_p2.diag[i])));
    }
    for (int r = 4; r < 8; r++) {
        for (int i = 0; i < 16; i++) s[i] = vsbox(vadd(s[i], vset(g_p2.ext[r][i])));
        vext(s);
    }
}
static void perm_scalar(uint32_t st[16]) {               // one state through the 8-lane code (lane 0)
    V s[16];
    for (int i = 0; i < 16; i++) s[i] = vset(st[i]);
    vperm(s);
    for (int i = 0; i < 16; i++) st[i] = (uint32_t)_mm256_extract_epi32(s[i], 0);
}

}  // namespace fastc

extern "C" {

// One matrix, height 2^log_h, `width` columns. data: column-major canonical words, or NULL to generate the synthetic
// trace of orc_fill_synthetic(seed) column by column (so a 16 GiB trace never has to exist on the host). Writes the root
// (8 canonical words). times[3] (optional): seconds of trace generation / conversion, LDE, hashing (leaves + tree), each
// a sum of barrier-separated parallel phases. Returns 0, or -1 on bad arguments.
int orc_fast_commit(const uint32_t* data, uint64_t seed, uint32_t log_h, uint64_t width, uint32_t log_blowup, int threads,
                    uint32_t root_out[8], double times[3]) {
    using namespace fastc;
    if (log_h + log_blowup > 27 || width == 0 || log_blowup > 4) return -1;
    init_constants();
    init_p2();
    if (threads < 1) threads = 1;
    const size_t n = (size_t)1 << log_h, B = (size_t)1 << log_blowup, m = n * B;
    const unsigned logm = log_h + log_blowup;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto secs = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double>(b - a).count();
    };
    double t_gen = 0, t_lde = 0, t_hash = 0;
    auto t0 = now();
    // tables: inverse stages; forward stages; per coset the scale (GENERATOR w_M^j)^i / N
    const uint32_t gn = to_m(orc::two_adic_generator(log_h).v);
    std::vector<uint32_t> tw_fwd, tw_inv, scale(m);
    if (log_h) {
        tw_fwd = stage_table(log_h, gn, threads);
        tw_inv = stage_table(log_h, mpow(gn, P - 2), threads);
    }
    {
        const uint32_t ninv = mpow(to_m((uint32_t)(n % P)), P - 2);
        const uint32_t wm = to_m(orc::two_adic_generator(logm).v), g = to_m(orc::GENERATOR);
        for (size_t j = 0; j < B; j++) {
            const uint32_t base = mmul(g, mpow(wm, j));
            uint32_t* dst = scale.data() + j * n;
            const size_t seg = 1 << 12;
            par_items((n + seg - 1) / seg, threads, [&](size_t q) {
                const size_t lo = q * seg, hi = std::min(n, lo + seg);
                uint32_t x = mmul(ninv, mpow(base, lo));
                for (size_t i = lo; i < hi; i++) { dst[i] = x; x = mmul(x, base); }
            });
        }
    }
    t_lde += secs(t0, now());

    // columns in batches of T (a multiple of 8 = one sponge block, except the ragged last batch)
    size_t T = ((size_t)threads + 7) / 8 * 8;
    if (T > width) T = width;
    std::vector<uint32_t> batch(T * m), work(std::min<size_t>(T, (size_t)threads) * n);
    const bool single_batch = T >= width;
    std::vector<uint32_t> state((single_batch && m >= 8) ? 0 : 16 * m, 0), digests(8 * m);
    for (size_t c0 = 0; c0 < width; c0 += T) {
        const size_t cw = std::min(T, width - c0);
        // (1) load / generate + convert to Montgomery form, (2) LDE: one column per thread, private workspace
        std::vector<double> gen_s(cw, 0.0);
        auto tl = now();
        std::atomic<size_t> slot_ctr{0};
        {
            std::atomic<size_t> next{0};
            const size_t nthr = std::min<size_t>((size_t)threads, cw);
            auto body = [&] {
                const size_t slot = slot_ctr.fetch_add(1);
                uint32_t* buf = work.data() + slot * n;   // evaluations -> coefficients (bit-reversed order)
                for (size_t k; (k = next.fetch_add(1)) < cw;) {
                    const size_t c = c0 + k;
                    auto g0 = now();
                    if (data) {
                        const uint32_t* src = data + c * n;
                        for (size_t r = 0; r < n; r++) buf[r] = to_m(src[r]);
                    } else {
                        for (size_t r = 0; r < n; r++) {
                            uint64_t x = seed ^ (((uint64_t)c << 32) | r);
                            x += 0x9E3779B97F4A7C15ull;
                            x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
                            x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
                            x ^= x >> 31;
                            buf[r] = to_m((uint32_t)(x % P));
                        }
                    }
                    gen_s[k] = secs(g0, now());
                    uint32_t* out = batch.data() + k * m;
                    if (log_h == 0) {
                        for (size_t i = 0; i < m; i++) out[i] = buf[0];
                        continue;
                    }
                    dif_ntt(buf, log_h, tw_inv.data());
                    for (size_t j = 0; j < B; j++) {
                        uint32_t* blk = out + (size_t)brev((uint32_t)j, log_blowup) * n;
                        bitrev_scale(blk, buf, scale.data() + j * n, log_h);
                        dif_ntt(blk, log_h, tw_fwd.data());
                    }
                }
            };
            std::vector<std::thread> th;
            for (size_t i = 0; i < nthr; i++) th.emplace_back(body);
            for (auto& x : th) x.join();
        }
        {
            // generation runs inside the same parallel phase: book its share of the phase's wall time separately
            double g = 0;
            for (double x : gen_s) g += x;
            const double wall = secs(tl, now()), share = g / std::max<double>(1.0, (double)std::min<size_t>((size_t)threads, cw));
            t_gen += std::min(wall, share);
            t_lde += wall - std::min(wall, share);
        }
        // (3) absorb the batch into the per-row sponge states, 8 rows per vector
        auto th0 = now();
        const bool first = c0 == 0, last = c0 + cw >= width;
        par_range(m / 8 ? m / 8 : 1, threads, [&](size_t lo, size_t hi) {
            if (m < 8) {   // tiny matrices: scalar lanes
                for (size_t r = 0; r < m; r++) {
                    uint32_t st[16];
                    for (int i = 0; i < 16; i++) st[i] = first ? 0 : state[(size_t)i * m + r];
                    for (size_t c = 0; c < cw; c += 8) {
                        for (size_t k = 0; k < 8 && c + k < cw; k++) st[k] = batch[(c + k) * m + r];
                        perm_scalar(st);
                    }
                    for (int i = 0; i < 16; i++) state[(size_t)i * m + r] = st[i];
                }
                return;
            }
            for (size_t v = lo; v < hi; v++) {
                const size_t r = v * 8;
                V s[16];
                for (int i = 0; i < 16; i++) s[i] = first ? _mm256_setzero_si256() : vload(&state[(size_t)i * m + r]);
                for (size_t c = 0; c < cw; c += 8) {
                    for (size_t k = 0; k < 8 && c + k < cw; k++) s[k] = vload(&batch[(c + k) * m + r]);
                    vperm(s);
                }
                if (last) { for (int i = 0; i < 8; i++) vstore(&digests[(size_t)i * m + r], s[i]); }
                else { for (int i = 0; i < 16; i++) vstore(&state[(size_t)i * m + r], s[i]); }
            }
        });
        if (m < 8 && last)
            for (int i = 0; i < 8; i++) for (size_t r = 0; r < m; r++) digests[(size_t)i * m + r] = state[(size_t)i * m + r];
        t_hash += secs(th0, now());
    }
    // (4) tree: planes of 8 words per layer, compress(left, right) = first 8 words of permute(left || right)
    auto tt = now();
    std::vector<uint32_t> cur = std::move(digests);
    size_t h = m;
    while (h > 1) {
        const size_t nh = h / 2;
        std::vector<uint32_t> nxt(8 * nh);
        if (nh >= 8) {
            const V idx = _mm256_setr_epi32(0, 2, 4, 6, 8, 10, 12, 14);
            par_range(nh / 8, threads, [&](size_t lo, size_t hi) {
                for (size_t v = lo; v < hi; v++) {
                    const size_t i = v * 8;
                    V s[16];
                    for (int w = 0; w < 8; w++) {
                        const int* base = (const int*)&cur[(size_t)w * h + 2 * i];
                        s[w] = _mm256_i32gather_epi32(base, idx, 4);
                        s[8 + w] = _mm256_i32gather_epi32(base + 1, idx, 4);
                    }
                    vperm(s);
                    for (int w = 0; w < 8; w++) vstore(&nxt[(size_t)w * nh + i], s[w]);
                }
            });
        } else {
            for (size_t i = 0; i < nh; i++) {
                uint32_t st[16];
                for (int w = 0; w < 8; w++) { st[w] = cur[(size_t)w * h + 2 * i]; st[8 + w] = cur[(size_t)w * h + 2 * i + 1]; }
                perm_scalar(st);
                for (int w = 0; w < 8; w++) nxt[(size_t)w * nh + i] = st[w];
            }
        }
        cur = std::move(nxt);
        h = nh;
    }
    for (int w = 0; w < 8; w++) root_out[w] = from_m(cur[w]);
    t_hash += secs(tt, now());
    if (times) { times[0] = t_gen; times[1] = t_lde; times[2] = t_hash; }
    return 0;
}

}  // extern "C"
