// Poseidon2-BabyBear width 16 on the host with AVX2: the Fiat-Shamir transcript of a proof is a sequential chain of a few
// thousand permutations on the CPU (host_challenger.hpp; the reference's challenger is host-side and sequential too), and
// with the opened values of a wide AIR it was the longest host-only stretch of a proof (KeccakVmAir shape: 6.9 ms of 81).
// Same permutation, same Montgomery words as poseidon2.cuh / HostPoseidon2::permute_scalar; the state is two vectors of
// eight lanes, the S-box of the full rounds and the diagonal of the partial rounds run eight lanes at a time.
// Compiled with -mavx2 (openvm_b200/build.py); callers ask b2_host_has_avx2() first.
#include <cstdint>
#if defined(__x86_64__)
#include <immintrin.h>
#include "poseidon2_constants.cuh"

namespace {
constexpr uint32_t P = 0x78000001u;
constexpr uint32_t P_NEG_INV = 0x77FFFFFFu;  // -p^-1 mod 2^32

alignas(32) const uint32_t EXT[8][16] = B2_P2_EXT_ALL;
const uint32_t PARTIAL[13] = B2_P2_PARTIAL;
alignas(32) const uint32_t DIAG[16] = B2_P2_DIAG;

inline __m256i p8() { return _mm256_set1_epi32((int)P); }
inline __m256i add_mod(__m256i a, __m256i b) {  // a, b < p: the sum fits 32 bits
    const __m256i s = _mm256_add_epi32(a, b);
    return _mm256_min_epu32(s, _mm256_sub_epi32(s, p8()));
}
// Montgomery product of eight lanes: even and odd lanes as 64-bit products, (t + (t * -p^-1 mod 2^32) * p) >> 32 in [0, 2p)
inline __m256i mul_mod(__m256i a, __m256i b) {
    const __m256i ninv = _mm256_set1_epi32((int)P_NEG_INV), p = p8();
    const __m256i te = _mm256_mul_epu32(a, b);
    const __m256i to = _mm256_mul_epu32(_mm256_srli_epi64(a, 32), _mm256_srli_epi64(b, 32));
    const __m256i ue = _mm256_mul_epu32(_mm256_mul_epu32(te, ninv), p);
    const __m256i uo = _mm256_mul_epu32(_mm256_mul_epu32(to, ninv), p);
    const __m256i se = _mm256_add_epi64(te, ue), so = _mm256_add_epi64(to, uo);
    const __m256i r = _mm256_blend_epi32(_mm256_srli_epi64(se, 32), so, 0xAA);
    return _mm256_min_epu32(r, _mm256_sub_epi32(r, p));
}
inline __m256i sbox(__m256i x, __m256i rc) {
    const __m256i y = add_mod(x, rc);
    const __m256i y2 = mul_mod(y, y), y3 = mul_mod(y2, y), y4 = mul_mod(y2, y2);
    return mul_mod(y3, y4);
}
inline uint32_t mul_scalar(uint32_t a, uint32_t b) {
    const uint64_t t = (uint64_t)a * b;
    const uint32_t m = (uint32_t)t * P_NEG_INV;
    const uint32_t r = (uint32_t)((t + (uint64_t)m * P) >> 32);
    return r >= P ? r - P : r;
}
// circ(2, 3, 1, 1) on each block of four lanes: out_i = (x_i + x_i+1 + x_i+2 + x_i+3) + x_i + 2 x_i+1
inline __m256i mat4(__m256i x) {
    const __m256i r1 = _mm256_shuffle_epi32(x, 0x39);            // x_i+1
    const __m256i t = add_mod(x, r1);                             // x_i + x_i+1
    const __m256i sum = add_mod(t, _mm256_shuffle_epi32(t, 0x4E));
    return add_mod(add_mod(sum, t), r1);
}
inline void external(__m256i& a, __m256i& b) {
    a = mat4(a);
    b = mat4(b);
    __m256i s = add_mod(a, b);                                    // blocks 0 + 2 | 1 + 3
    s = add_mod(s, _mm256_permute2x128_si256(s, s, 0x01));        // all four blocks, per lane of the block
    a = add_mod(a, s);
    b = add_mod(b, s);
}
}  // namespace

extern "C" int b2_host_has_avx2() { return __builtin_cpu_supports("avx2") ? 1 : 0; }

extern "C" void b2_poseidon2_permute_avx2(uint32_t* st) {
    __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(st));
    __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(st + 8));
    const __m256i da = _mm256_load_si256(reinterpret_cast<const __m256i*>(DIAG));
    const __m256i db = _mm256_load_si256(reinterpret_cast<const __m256i*>(DIAG + 8));
    external(a, b);
    for (int r = 0; r < 4; r++) {
        a = sbox(a, _mm256_load_si256(reinterpret_cast<const __m256i*>(EXT[r])));
        b = sbox(b, _mm256_load_si256(reinterpret_cast<const __m256i*>(EXT[r] + 8)));
        external(a, b);
    }
    for (int r = 0; r < 13; r++) {
        // S-box on lane 0 only
        uint32_t y = (uint32_t)_mm256_cvtsi256_si32(a) + PARTIAL[r];
        y = y >= P ? y - P : y;
        const uint32_t y2 = mul_scalar(y, y), y3 = mul_scalar(y2, y), y4 = mul_scalar(y2, y2);
        a = _mm256_blend_epi32(a, _mm256_castsi128_si256(_mm_cvtsi32_si128((int)mul_scalar(y3, y4))), 0x01);
        // s_i <- sum(s) + d_i s_i
        __m256i t = add_mod(a, b);
        t = add_mod(t, _mm256_permute2x128_si256(t, t, 0x01));
        t = add_mod(t, _mm256_shuffle_epi32(t, 0x4E));
        t = add_mod(t, _mm256_shuffle_epi32(t, 0xB1));
        a = add_mod(t, mul_mod(a, da));
        b = add_mod(t, mul_mod(b, db));
    }
    for (int r = 4; r < 8; r++) {
        a = sbox(a, _mm256_load_si256(reinterpret_cast<const __m256i*>(EXT[r])));
        b = sbox(b, _mm256_load_si256(reinterpret_cast<const __m256i*>(EXT[r] + 8)));
        external(a, b);
    }
    _mm256_storeu_si256(reinterpret_cast<__m256i*>(st), a);
    _mm256_storeu_si256(reinterpret_cast<__m256i*>(st + 8), b);
}
#else
extern "C" int b2_host_has_avx2() { return 0; }
extern "C" void b2_poseidon2_permute_avx2(uint32_t*) {}
#endif
