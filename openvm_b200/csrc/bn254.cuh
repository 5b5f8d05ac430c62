// BN254 scalar field and Poseidon2 width 3 over it: the hash of the reference's root configuration
// (BabyBearPoseidon2RootConfig, crates/continuations/src/lib.rs:17; crates/sdk/src/prover/root.rs), whose proofs the
// halo2 wrapper verifies with BN254-native hashing. In-tree twins this follows:
//   permutation (round structure, circ(2,1,1), diag [1,1,2] + J, x^5)  extensions/native/compiler/src/constraints/halo2/poseidon2_perm.rs:52-215
//   parameters (8 full + 56 partial rounds, zkhash RC3)                extensions/native/compiler/src/constraints/halo2/compiler.rs:39-85
//   sponge over packed BabyBear words, 2-to-1 compression              extensions/native/recursion/src/outer_poseidon2.rs:26-63
//   reduce_32 / split_32                                               extensions/native/recursion/src/utils.rs:42-58,
//                                                                      extensions/native/compiler/src/constraints/halo2/compiler.rs:603-632
// Host and device share this code (the verifier and the transcript run it on the host). Elements are eight 32-bit limbs,
// least significant first, in Montgomery form (R = 2^256); digests leave as canonical limbs.
#pragma once
#include "field.cuh"
#include "bn254_constants.cuh"

namespace b2 {

#if defined(__CUDACC__)
#define B2_BN_CALL __host__ __device__ __noinline__   // one copy of the 8 x 8 limb product per kernel (instruction cache)
static __constant__ uint32_t BN_EXT_D[8][3][8] = B2_BN_EXT_ALL;
static __constant__ uint32_t BN_PARTIAL_D[56][8] = B2_BN_PARTIAL;
#else
#define B2_BN_CALL inline
#endif
static const uint32_t BN_EXT_H[8][3][8] = B2_BN_EXT_ALL;
static const uint32_t BN_PARTIAL_H[56][8] = B2_BN_PARTIAL;

struct Fr {
    uint32_t l[8];
};

B2_HD Fr fr_zero() {
    Fr z;
#pragma unroll
    for (int i = 0; i < 8; i++) z.l[i] = 0;
    return z;
}
// t < 2 r  ->  t mod r
B2_HD void fr_cond_sub(Fr& t) {
    constexpr uint32_t N[8] = B2_BN_MODULUS;
    uint32_t d[8];
    uint64_t borrow = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint64_t x = (uint64_t)t.l[i] - N[i] - borrow;
        d[i] = (uint32_t)x;
        borrow = (x >> 32) & 1;
    }
    if (!borrow) {
#pragma unroll
        for (int i = 0; i < 8; i++) t.l[i] = d[i];
    }
}
B2_HD Fr fr_add(const Fr& a, const Fr& b) {  // a, b < r < 2^254: no carry out of the top limb
    Fr s;
    uint64_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        c += (uint64_t)a.l[i] + b.l[i];
        s.l[i] = (uint32_t)c;
        c >>= 32;
    }
    fr_cond_sub(s);
    return s;
}
// Montgomery product, operand scanning with interleaved reduction (CIOS): a b 2^-256 mod r for a b < r 2^256
static B2_BN_CALL Fr fr_mul(Fr a, Fr b) {
    constexpr uint32_t N[8] = B2_BN_MODULUS;
    uint32_t t[9];
#pragma unroll
    for (int i = 0; i < 9; i++) t[i] = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint64_t c = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint64_t x = (uint64_t)a.l[j] * b.l[i] + t[j] + c;
            t[j] = (uint32_t)x;
            c = x >> 32;
        }
        const uint64_t top = (uint64_t)t[8] + c;   // a b < r 2^256 keeps the running value below 2^288
        const uint32_t m = t[0] * B2_BN_NEG_INV32;
        c = ((uint64_t)m * N[0] + t[0]) >> 32;
#pragma unroll
        for (int j = 1; j < 8; j++) {
            const uint64_t x = (uint64_t)m * N[j] + t[j] + c;
            t[j - 1] = (uint32_t)x;
            c = x >> 32;
        }
        const uint64_t y = (uint64_t)(uint32_t)top + c;
        t[7] = (uint32_t)y;
        t[8] = (uint32_t)(top >> 32) + (uint32_t)(y >> 32);
    }
    Fr r;
#pragma unroll
    for (int i = 0; i < 8; i++) r.l[i] = t[i];
    fr_cond_sub(r);
    return r;
}
// any 256-bit integer (canonical limbs, not necessarily below r) -> Montgomery form of its residue
B2_HD Fr fr_from_words(const uint32_t (&w)[8]) {
    constexpr uint32_t R2[8] = B2_BN_R2;
    Fr a, b;
#pragma unroll
    for (int i = 0; i < 8; i++) { a.l[i] = w[i]; b.l[i] = R2[i]; }
    return fr_mul(a, b);
}
B2_HD void fr_to_words(const Fr& a, uint32_t (&w)[8]) {
    Fr one = fr_zero();
    one.l[0] = 1;
    const Fr c = fr_mul(a, one);
#pragma unroll
    for (int i = 0; i < 8; i++) w[i] = c.l[i];
}

B2_HD Fr bn_round_constant_ext(int r, int i) {
    Fr c;
#pragma unroll
    for (int k = 0; k < 8; k++) {
#if defined(__CUDA_ARCH__)
        c.l[k] = BN_EXT_D[r][i][k];
#else
        c.l[k] = BN_EXT_H[r][i][k];
#endif
    }
    return c;
}
B2_HD Fr bn_round_constant_partial(int r) {
    Fr c;
#pragma unroll
    for (int k = 0; k < 8; k++) {
#if defined(__CUDA_ARCH__)
        c.l[k] = BN_PARTIAL_D[r][k];
#else
        c.l[k] = BN_PARTIAL_H[r][k];
#endif
    }
    return c;
}
B2_HD Fr bn_sbox(const Fr& x) {
    const Fr x2 = fr_mul(x, x), x4 = fr_mul(x2, x2);
    return fr_mul(x, x4);
}
B2_HD void bn_external(Fr (&s)[3]) {  // circ(2, 1, 1)
    const Fr t = fr_add(fr_add(s[0], s[1]), s[2]);
    s[0] = fr_add(s[0], t);
    s[1] = fr_add(s[1], t);
    s[2] = fr_add(s[2], t);
}
B2_HD void bn_internal(Fr (&s)[3]) {  // J + diag(1, 1, 2)
    const Fr t = fr_add(fr_add(s[0], s[1]), s[2]);
    s[0] = fr_add(s[0], t);
    s[1] = fr_add(s[1], t);
    s[2] = fr_add(fr_add(s[2], s[2]), t);
}
B2_HD void poseidon2_bn254_permute(Fr (&s)[3]) {
    bn_external(s);
#pragma unroll 1
    for (int r = 0; r < 8; r++) {
        if (r == 4) {
#pragma unroll 1
            for (int q = 0; q < 56; q++) {
                s[0] = bn_sbox(fr_add(s[0], bn_round_constant_partial(q)));
                bn_internal(s);
            }
        }
        s[0] = bn_sbox(fr_add(s[0], bn_round_constant_ext(r, 0)));
        s[1] = bn_sbox(fr_add(s[1], bn_round_constant_ext(r, 1)));
        s[2] = bn_sbox(fr_add(s[2], bn_round_constant_ext(r, 2)));
        bn_external(s);
    }
}

// ---- BabyBear words in and out of an element ---------------------------------------------------------------------------
// reduce_32: sum_k v_k 2^(32 k) over up to eight canonical BabyBear values (the sponge packs 8 = 254 / 31 per element,
// the challenger 3 = 254 / 64); split_32: the first n 64-bit limbs of the canonical integer, each reduced mod p.
B2_HD Fr bn_reduce_32(const uint32_t (&canonical)[8]) { return fr_from_words(canonical); }
B2_HD void bn_split_32(const Fr& x, uint32_t (&felts)[3]) {
    uint32_t w[8];
    fr_to_words(x, w);
#pragma unroll
    for (int j = 0; j < 3; j++) felts[j] = (uint32_t)((((uint64_t)w[2 * j + 1] << 32) | w[2 * j]) % P);
}
// TruncatedPermutation<Perm, 2, 1, 3> on canonical digests
B2_HD void bn_compress(const uint32_t (&l)[8], const uint32_t (&r)[8], uint32_t (&out)[8]) {
    Fr s[3] = {fr_from_words(l), fr_from_words(r), fr_zero()};
    poseidon2_bn254_permute(s);
    fr_to_words(s[0], out);
}

}  // namespace b2
