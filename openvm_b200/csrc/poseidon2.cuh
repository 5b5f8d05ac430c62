// Poseidon2-BabyBear width 16 (x^7; 4 + 13 + 4 rounds), one permutation per thread, state in registers.
//
// Replaces (reference, external): p3-baby-bear `Poseidon2BabyBear<16>` with zkhash RC16 constants as
// configured at extensions/native/recursion/src/hints.rs:35 and
// crates/circuits/poseidon2-air/src/{lib.rs:37-44, babybear.rs:20-64, permute.rs:104-146};
// layer structure as in cuda/include/poseidon2-air/linearlayer.cuh:13-94.
// Also the sponge / compression built on it:
//   PaddingFreeSponge<Perm,16,8,8>      hints.rs:36, extensions/native/circuit/src/poseidon2/README.md:9-17
//   TruncatedPermutation<Perm,2,8,16>   hints.rs:37-38, crates/vm/src/arch/hasher/poseidon2.rs:33-43
#pragma once
#include "field.cuh"
#include "poseidon2_constants.cuh"

namespace b2 {

#if defined(__CUDACC__)
// round constants: rows 0-3 initial full rounds, rows 4-7 terminal full rounds (one table so a single looped
// code body serves all eight full rounds; indexed by the loop counter -> LDC with a register index)
static __constant__ uint32_t P2C_EXT[8][16] = B2_P2_EXT_ALL;
static __constant__ uint32_t P2C_PARTIAL[13] = B2_P2_PARTIAL;

// Signed Montgomery product (a, b in (-p, p) as int32 -> a*b*2^-32 in (-p, p)): r = hi(t) - hi(m*p) with
// m = lo(t) * p^-1 needs no final correction because |t| < 2^31 p keeps both high words within p/2. A chain of
// products therefore costs 3 multiplies + 1 subtract each instead of 3 + 2 (the unsigned form corrects into [0, p)
// after every product). The fmaheavy pipe carries the multiplies either way; this trims the ALU side so that ptxas
// can move more of the additions it currently issues as IMAD.IADD off the multiply pipe (profiles/r1c: fmaheavy 87 %).
B2_D int32_t smont(int32_t a, int32_t b) {
    const int64_t t = (int64_t)a * b;
    const uint32_t m = (uint32_t)t * P_INV;
    const int32_t u = __mulhi((int32_t)m, (int32_t)P);
    return (int32_t)(t >> 32) - u;
}
// x -> (x + rc)^7 for x, rc in [0, p): the constant is added as rc - p, which lands in (-p, p) with one 3-input add,
// the four products run signed, and a single add-min brings the result back to [0, p)
B2_D Fp sbox7_rc(Fp x, uint32_t rc) {
    const int32_t y = (int32_t)(x.v + rc - P);
    const int32_t y2 = smont(y, y);
    const int32_t y3 = smont(y2, y);
    const int32_t y4 = smont(y2, y2);
    const uint32_t y7 = (uint32_t)smont(y3, y4);
    const uint32_t c = y7 + P;
    return Fp::raw(c < y7 ? c : y7);
}

// x * 2^-K mod p without a Montgomery product: p = 15 * 2^27 + 1 gives 2^-27 = -15, so 2^-K = -15 * 2^(27-K) and
//   x * 2^-K = (x >> K) - 15 * 2^(27-K) * (x mod 2^K)      (second term < 15 * 2^27 < p)
// one 2-cycle IMAD instead of a 10-cycle product on the multiply pipe. NEG gives -(x * 2^-K).
template <int K, bool NEG>
B2_D Fp mul_2exp_neg(Fp x) {
    static_assert(K >= 1 && K <= 27, "K");
    const uint32_t lo = x.v & ((1u << K) - 1), hi = x.v >> K;
    const uint32_t t = lo * (15u << (27 - K));
    const uint32_t r = NEG ? t - hi : hi - t;
    const uint32_t c = r + P;
    return Fp::raw(c < r ? c : r);
}

// [[2,3,1,1],[1,2,3,1],[1,1,2,3],[3,1,1,2]] on four lanes: 7 additions + 2 doublings
B2_D void mat4(Fp& x0, Fp& x1, Fp& x2, Fp& x3) {
    Fp t01 = x0 + x1;
    Fp t23 = x2 + x3;
    Fp t0123 = t01 + t23;
    Fp t01123 = t0123 + x1;
    Fp t01233 = t0123 + x3;
    Fp n3 = t01233 + fp_double(x0);  // 3x0 + x1 + x2 + 2x3
    Fp n1 = t01123 + fp_double(x2);  // x0 + 2x1 + 3x2 + x3
    Fp n0 = t01123 + t01;            // 2x0 + 3x1 + x2 + x3
    Fp n2 = t01233 + t23;            // x0 + x1 + 2x2 + 3x3
    x0 = n0; x1 = n1; x2 = n2; x3 = n3;
}

// MDS-light external layer: circ(2*M4, M4, M4, M4)
B2_D void p2_external(Fp (&s)[16]) {
#pragma unroll
    for (int b = 0; b < 16; b += 4) mat4(s[b], s[b + 1], s[b + 2], s[b + 3]);
    Fp sums[4];
#pragma unroll
    for (int k = 0; k < 4; k++) sums[k] = (s[k] + s[4 + k]) + (s[8 + k] + s[12 + k]);
#pragma unroll
    for (int i = 0; i < 16; i++) s[i] += sums[i & 3];
}

// halve: x/2 mod p without a Montgomery product
B2_D Fp fp_halve(Fp a) {
    uint32_t odd = a.v & 1u;
    return Fp::raw((a.v >> 1) + odd * ((P + 1) >> 1));
}

// internal layer: s_i <- sum(s) + d_i * s_i   with d = (-2, 1, 2, 1/2, 3, 4, -1/2, -3, -4, 2^-8, 1/4, 1/8, 2^-27, -2^-8, -1/16, -2^-27)
// No Montgomery product at all: small-integer and 1/2 multiples are additions/shifts, the 2^-k entries use the
// shape of p (mul_2exp_neg); the fmaheavy pipe, which carries the S-box products, is the busier one.
B2_D void p2_internal(Fp (&s)[16]) {
    Fp part = ((s[1] + s[2]) + (s[3] + s[4])) + ((s[5] + s[6]) + (s[7] + s[8])) +
              (((s[9] + s[10]) + (s[11] + s[12])) + ((s[13] + s[14]) + s[15]));
    Fp full = part + s[0];
    s[0] = part - s[0];
    s[1] = full + s[1];
    s[2] = full + fp_double(s[2]);
    {
        Fp h3 = fp_halve(s[3]);
        Fp d4 = fp_double(s[4]), d5 = fp_double(s[5]);
        Fp h6 = fp_halve(s[6]);
        Fp d7 = fp_double(s[7]), d8 = fp_double(s[8]);
        s[3] = full + h3;
        s[4] = full + (d4 + s[4]);
        s[5] = full + fp_double(d5);
        s[6] = full - h6;
        s[7] = full - (d7 + s[7]);
        s[8] = full - fp_double(d8);
    }
    s[9] = full + mul_2exp_neg<8, false>(s[9]);
    s[10] = full + mul_2exp_neg<2, false>(s[10]);
    s[11] = full + mul_2exp_neg<3, false>(s[11]);
    s[12] = full + mul_2exp_neg<27, false>(s[12]);
    s[13] = full + mul_2exp_neg<8, true>(s[13]);
    s[14] = full + mul_2exp_neg<4, true>(s[14]);
    s[15] = full + mul_2exp_neg<27, true>(s[15]);
}

// The whole permutation as  E S0 E S1 E S2 E S3 E [P x 13] S4 E S5 E S6 E S7 E  with one code body per layer type
// (E external layer, S_r add-constants + x^7 on all lanes, P partial round). Loops are kept rolled on purpose:
// fully unrolled, one permutation is ~100 KB of SASS, which thrashes the instruction cache (ncu: stall
// no_instruction dominant, profiles/r1a_ncu_full_summary.md); rolled it is ~9 KB.
B2_D void poseidon2_permute(Fp (&s)[16]) {
#pragma unroll 1
    for (int r = 0; r < 9; r++) {
        p2_external(s);
        if (r == 4) {
#pragma unroll 1
            for (int q = 0; q < 13; q++) {
                s[0] = sbox7_rc(s[0], P2C_PARTIAL[q]);
                p2_internal(s);
            }
        }
        if (r < 8) {
#pragma unroll
            for (int i = 0; i < 16; i++) s[i] = sbox7_rc(s[i], P2C_EXT[r][i]);
        }
    }
}
#endif  // __CUDACC__

}  // namespace b2
