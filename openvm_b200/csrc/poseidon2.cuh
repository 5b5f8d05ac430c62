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
static __constant__ uint32_t P2C_DIAG[16] = B2_P2_DIAG;

// Montgomery product whose  m = lo * (-p^-1)  step runs on the ALU pipe as shifts/adds instead of a 4-cycle IMAD
// on fmaheavy: -p^-1 = 2^31 - 2^27 - 1 (mod 2^32). Used for a fraction of the S-box products to balance the two
// pipes (ncu: fmaheavy 87 % busy, alu 53 %, profiles/r1c).
B2_D Fp mul_alu(Fp a, Fp b) {
    const uint64_t t = (uint64_t)a.v * b.v;
    const uint32_t lo = (uint32_t)t;
    uint32_t m;  // inline PTX so the compiler does not fold the shifts back into a multiply by 0x77ffffff
    asm("{\n\t.reg .u32 a, b;\n\tshl.b32 a, %1, 31;\n\tshl.b32 b, %1, 27;\n\tsub.u32 a, a, b;\n\tsub.u32 %0, a, %1;\n\t}"
        : "=r"(m) : "r"(lo));
    const uint64_t u = t + (uint64_t)m * P;
    const uint32_t r = (uint32_t)(u >> 32);
    const uint32_t s = r - P;
    return Fp::raw(s < r ? s : r);
}
template <int MASK>
B2_D Fp sbox7(Fp x) {
    Fp x2 = (MASK & 1) ? mul_alu(x, x) : x * x;
    Fp x3 = (MASK & 2) ? mul_alu(x2, x) : x2 * x;
    Fp x4 = (MASK & 4) ? mul_alu(x2, x2) : x2 * x2;
    return (MASK & 8) ? mul_alu(x3, x4) : x3 * x4;
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
// The small-integer and 1/2 multiples are done with additions/shifts on the ALU pipe (the fmaheavy pipe, which
// carries the Montgomery products, is the busier one); the negative-power-of-two entries use Montgomery products.
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
#pragma unroll
    for (int i = 9; i < 16; i++) s[i] = full + s[i] * Fp::raw(P2C_DIAG[i]);
}

// The whole permutation as  E S0 E S1 E S2 E S3 E [P x 13] S4 E S5 E S6 E S7 E  with one code body per layer type
// (E external layer, S_r add-constants + x^7 on all lanes, P partial round). Loops are kept rolled on purpose:
// fully unrolled, one permutation is ~100 KB of SASS, which thrashes the instruction cache (ncu: stall
// no_instruction dominant, profiles/r1a_ncu_full_summary.md); rolled it is ~9 KB.
#ifndef B2_P2_FULL_MASK
#define B2_P2_FULL_MASK 0
#endif
#ifndef B2_P2_PARTIAL_MASK
#define B2_P2_PARTIAL_MASK 0
#endif
template <int FULL_MASK = B2_P2_FULL_MASK, int PARTIAL_MASK = B2_P2_PARTIAL_MASK>
B2_D void poseidon2_permute(Fp (&s)[16]) {
#pragma unroll 1
    for (int r = 0; r < 9; r++) {
        p2_external(s);
        if (r == 4) {
#pragma unroll 1
            for (int q = 0; q < 13; q++) {
                s[0] = sbox7<PARTIAL_MASK>(s[0] + Fp::raw(P2C_PARTIAL[q]));
                p2_internal(s);
            }
        }
        if (r < 8) {
#pragma unroll
            for (int i = 0; i < 16; i++) s[i] = sbox7<FULL_MASK>(s[i] + Fp::raw(P2C_EXT[r][i]));
        }
    }
}
#endif  // __CUDACC__

}  // namespace b2
