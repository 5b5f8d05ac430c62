// BabyBear arithmetic on the FP64 pipe: field elements as exact integers held in doubles.
//
// A B200 issues a DFMA per 2 cycles per SM sub-partition -- the rate of a 32-bit IMAD -- on a pipe the integer code
// never touches. Two uses (profiles/r1g_fp64_pipe.md):
//   * products with an immediate reduction (dmod / dmodc): the Poseidon2 permutation of poseidon2_f64.cuh, run next to
//     the integer one;
//   * dot products with NO reduction inside the loop: a 31-bit value times a 16-bit half of the other factor is below
//     2^47, so 64 such products accumulate exactly in one double by plain DFMA (reduced openings, prover.cu).
#pragma once
#include "field.cuh"

namespace b2 {

constexpr uint32_t C_RINV = c_inv(R_MOD_P);  // 2^-32 mod p: Montgomery word -> canonical value

#if defined(__CUDACC__)
constexpr double D_P = 2013265921.0;
constexpr double D_INVP = 1.0 / 2013265921.0;
constexpr double D_MAGIC = 6755399441055744.0;     // 1.5 * 2^52: x + MAGIC - MAGIC = round to nearest integer, |x| < 2^51
constexpr double D_P_HIGH = 2013265920.0;          // p - 1 = 15 * 2^27: q * (p - 1) is exact for |q| < 2^49
constexpr double D_2_52 = 4503599627370496.0;
constexpr double D_RINV = (double)C_RINV;
constexpr double D_RINV_OVER_P = (double)C_RINV / 2013265921.0;
constexpr double D_R = (double)R_MOD_P;
constexpr double D_R_OVER_P = (double)R_MOD_P / 2013265921.0;

B2_D double dmod(double a, double b) {
    const double h = a * b;
    const double l = fma(a, b, -h);
    const double q = fma(h, D_INVP, D_MAGIC) - D_MAGIC;
    return fma(-q, D_P, h) + l;
}
// x * c mod p for a constant c < p with c_over_p = c / p;  |x| < 2^40
B2_D double dmodc(double x, double c, double c_over_p) {
    const double q = fma(x, c_over_p, D_MAGIC) - D_MAGIC;
    return fma(x, c, -(q * D_P_HIGH)) - q;
}
B2_D double dred(double x) {  // |x| < 2^51 -> |r| <= p/2 + 1
    const double q = fma(x, D_INVP, D_MAGIC) - D_MAGIC;
    return fma(-q, D_P, x);
}
// exact u32 -> double without a conversion instruction (2^52 + x has x in its low mantissa word)
B2_D double f64_from_u32(uint32_t x) { return __hiloint2double(0x43300000, (int)x) - D_2_52; }
// integer-valued double in (-p, p) -> u32 in [0, p)
B2_D uint32_t f64_to_u32_modp(double r) {
    const double t = (r < 0.0 ? r + D_P : r) + D_2_52;
    return (uint32_t)__double2loint(t);
}
// Montgomery word of the integer code path <-> canonical double
B2_D double f64_from_monty(uint32_t m) {
    const double x = __hiloint2double(0x43300000, (int)m) - D_2_52;   // exact u32 -> double without a conversion instruction
    return dmodc(x, D_RINV, D_RINV_OVER_P);
}
B2_D uint32_t f64_to_monty(double x) {  // |x| < 2^40
    const double r = dmodc(x, D_R, D_R_OVER_P);
    const double t = (r < 0.0 ? r + D_P : r) + D_2_52;
    return (uint32_t)__double2loint(t);
}
#endif  // __CUDACC__

}  // namespace b2
