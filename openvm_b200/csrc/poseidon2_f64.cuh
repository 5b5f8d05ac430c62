// Poseidon2-BabyBear width 16 on the FP64 pipe (same permutation as poseidon2.cuh, bit-identical results).
//
// Why: the integer permutation keeps the fmaheavy pipe 91 % busy (profiles/r1e_ncu_hash_summary.md) while the FP64
// pipe of a B200 idles. A permutation whose field elements are integers held in doubles runs entirely on DFMA / DADD /
// DMUL (3.3 G permutations/s alone, 97 % of that pipe's rate); with some warps of a block on this code path and the
// others on the integer one, both pipes work and the leaf hash gains ~12 % (tools/p2_fp64.cu, profiles/r1g_fp64_pipe.md;
// the limit is the shared issue port, not either pipe).
//
// Arithmetic: canonical domain (not Montgomery), values are exact integers |x| < 2^53 in a double.
//   product  (h, l) = error-free a*b;  q = round(h / p) via the 1.5*2^52 constant;  r = (h - q p) + l,  |r| <= 0.6 p,
//            exact for |a b| < 2^82 (q < 2^51); 6 operations
//   product by a constant c (c / p precomputed): q = round(x c / p);  r = fma(x, c, -q * 15*2^27) - q  (p = 15*2^27 + 1),
//            exact for |x| < 2^40; 5 operations
//   additions are plain DADD / DFMA with no correction; magnitudes are bounded per layer (comments below) and folded with
//   the 3-operation reduction dred where they would outgrow the product's domain.
#pragma once
#include "poseidon2.cuh"
#include "field_f64.cuh"

namespace b2 {

struct P2F64Tables {
    double ext[8][16];     // canonical round constants, full rounds
    double partial[13];    // canonical round constants, partial rounds
    double diag[16];       // canonical internal diagonal
    double diag_over_p[16];
};
constexpr P2F64Tables make_p2_f64_tables() {
    P2F64Tables t{};
    constexpr uint32_t ext[8][16] = B2_P2_EXT_ALL;
    constexpr uint32_t partial[13] = B2_P2_PARTIAL;
    constexpr uint32_t diag[16] = B2_P2_DIAG;
    for (int r = 0; r < 8; r++)
        for (int i = 0; i < 16; i++) t.ext[r][i] = (double)c_mul(ext[r][i], C_RINV);
    for (int q = 0; q < 13; q++) t.partial[q] = (double)c_mul(partial[q], C_RINV);
    for (int i = 0; i < 16; i++) {
        t.diag[i] = (double)c_mul(diag[i], C_RINV);
        t.diag_over_p[i] = t.diag[i] / 2013265921.0;
    }
    return t;
}

#if defined(__CUDACC__)
static __constant__ P2F64Tables P2D = make_p2_f64_tables();

B2_D double dsbox(double x, double rc) {  // |x| < 2^40.5 -> |r| <= p/2 + 1
    const double y = x + rc;
    const double y2 = dmod(y, y), y3 = dmod(y2, y), y4 = dmod(y2, y2);
    return dmod(y3, y4);
}
// |x_i| <= B  ->  |out_i| <= 35 B   (circ(2 M4, M4, M4, M4): non-negative entries, row sums 35)
B2_D void dexternal(double (&s)[16]) {
#pragma unroll
    for (int b = 0; b < 16; b += 4) {
        const double x0 = s[b], x1 = s[b + 1], x2 = s[b + 2], x3 = s[b + 3];
        const double t01 = x0 + x1, t23 = x2 + x3, t0123 = t01 + t23;
        const double t01123 = t0123 + x1, t01233 = t0123 + x3;
        s[b + 3] = fma(2.0, x0, t01233);
        s[b + 1] = fma(2.0, x2, t01123);
        s[b] = t01123 + t01;
        s[b + 2] = t01233 + t23;
    }
    double sums[4];
#pragma unroll
    for (int k = 0; k < 4; k++) sums[k] = (s[k] + s[4 + k]) + (s[8 + k] + s[12 + k]);
#pragma unroll
    for (int i = 0; i < 16; i++) s[i] += sums[i & 3];
}
// s_i <- sum(s) + d_i s_i,  d = (-2, 1, 2, 1/2, 3, 4, -1/2, -3, -4, 2^-8, 1/4, 1/8, 2^-27, -2^-8, -1/16, -2^-27).
// The sum is reduced every round (|full| <= p/2). Lanes with a fractional d_i go through a product (|s_i| <= p after it);
// the six small-integer lanes 1, 2, 4, 5, 7, 8 grow by at most 4 x + p/2 per round and are folded when `fold` is set
// (rounds 0, 3, 6, 9, 12: at most 42.5 p in between, p/2 after the last round); lane 0 is renewed by its S-box.
B2_D void dinternal(double (&s)[16], bool fold) {
    const double part = ((s[1] + s[2]) + (s[3] + s[4])) + ((s[5] + s[6]) + (s[7] + s[8])) +
                        (((s[9] + s[10]) + (s[11] + s[12])) + ((s[13] + s[14]) + s[15]));
    const double full = dred(part + s[0]);
    s[0] = fma(-2.0, s[0], full);
    s[1] = full + s[1];
    s[2] = fma(2.0, s[2], full);
    s[4] = fma(3.0, s[4], full);
    s[5] = fma(4.0, s[5], full);
    s[7] = fma(-3.0, s[7], full);
    s[8] = fma(-4.0, s[8], full);
    if (fold) {
        s[1] = dred(s[1]); s[2] = dred(s[2]); s[4] = dred(s[4]);
        s[5] = dred(s[5]); s[7] = dred(s[7]); s[8] = dred(s[8]);
    }
    s[3] = full + dmodc(s[3], P2D.diag[3], P2D.diag_over_p[3]);
    s[6] = full + dmodc(s[6], P2D.diag[6], P2D.diag_over_p[6]);
#pragma unroll
    for (int i = 9; i < 16; i++) s[i] = full + dmodc(s[i], P2D.diag[i], P2D.diag_over_p[i]);
}
// in: |s_i| <= 17.5 p (the output bound, so sponge chains need no reduction between permutations);
// out: |s_i| <= 17.5 p, congruent to the permutation's output; dred() brings a word to |x| <= p/2.
//   first external layer <= 612.5 p < 2^40.3 -> S-box -> p/2 -> external 17.5 p -> ... ; the partial rounds start from
//   17.5 p and leave every lane <= 1.5 p, the external layer after them <= 52.5 p.
B2_D void poseidon2_permute_f64(double (&s)[16]) {
#pragma unroll 1
    for (int r = 0; r < 9; r++) {
        dexternal(s);
        if (r == 4) {
#pragma unroll 1
            for (int q = 0; q < 13; q++) {
                s[0] = dsbox(s[0], P2D.partial[q]);
                dinternal(s, q % 3 == 0);
            }
        }
        if (r < 8) {
#pragma unroll
            for (int i = 0; i < 16; i++) s[i] = dsbox(s[i], P2D.ext[r][i]);
        }
    }
}
#endif  // __CUDACC__

}  // namespace b2
