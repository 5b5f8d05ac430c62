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
static __constant__ uint32_t P2C_EXT_INIT[4][16] = B2_P2_EXT_INIT;
static __constant__ uint32_t P2C_PARTIAL[13] = B2_P2_PARTIAL;
static __constant__ uint32_t P2C_EXT_TERM[4][16] = B2_P2_EXT_TERM;
static __constant__ uint32_t P2C_DIAG[16] = B2_P2_DIAG;

B2_D Fp sbox7(Fp x) {
    Fp x2 = x * x;
    Fp x3 = x2 * x;
    Fp x4 = x2 * x2;
    return x3 * x4;
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

// internal layer: s_i <- sum(s) + d_i * s_i   with d = (-2, 1, 2, 1/2, 3, 4, -1/2, -3, -4, 2^-8, 1/4, 1/8, 2^-27, -2^-8, -1/16, -2^-27)
B2_D void p2_internal(Fp (&s)[16]) {
    Fp part = ((s[1] + s[2]) + (s[3] + s[4])) + ((s[5] + s[6]) + (s[7] + s[8])) +
              (((s[9] + s[10]) + (s[11] + s[12])) + ((s[13] + s[14]) + s[15]));
    Fp full = part + s[0];
    s[0] = part - s[0];
    s[1] = full + s[1];
    s[2] = full + fp_double(s[2]);
#pragma unroll
    for (int i = 3; i < 16; i++) s[i] = full + s[i] * Fp::raw(P2C_DIAG[i]);
}

B2_D void poseidon2_permute(Fp (&s)[16]) {
    p2_external(s);
#pragma unroll
    for (int r = 0; r < 4; r++) {
#pragma unroll
        for (int i = 0; i < 16; i++) s[i] = sbox7(s[i] + Fp::raw(P2C_EXT_INIT[r][i]));
        p2_external(s);
    }
#pragma unroll
    for (int r = 0; r < 13; r++) {
        s[0] = sbox7(s[0] + Fp::raw(P2C_PARTIAL[r]));
        p2_internal(s);
    }
#pragma unroll
    for (int r = 0; r < 4; r++) {
#pragma unroll
        for (int i = 0; i < 16; i++) s[i] = sbox7(s[i] + Fp::raw(P2C_EXT_TERM[r][i]));
        p2_external(s);
    }
}
#endif  // __CUDACC__

}  // namespace b2
