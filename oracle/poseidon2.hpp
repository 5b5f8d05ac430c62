// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// Poseidon2-BabyBear width 16, as used by the reference's BabyBearPoseidon2Config
// (extensions/native/recursion/src/hints.rs:33-51).
//   structure: crates/circuits/poseidon2-air/src/lib.rs:37-44 (width 16, 4+13+4 rounds, x^7),
//              crates/circuits/poseidon2-air/src/permute.rs:104-146 (initial external layer, then rounds),
//              crates/circuits/poseidon2-air/cuda/include/poseidon2-air/linearlayer.cuh:13-94
//              (4x4 block [[2,3,1,1],[1,2,3,1],[1,1,2,3],[3,1,1,2]], circulant-light outer mix,
//               internal layer s_i <- sum + d_i*s_i with d_0=-2, d_1=1, d_2=2),
//              crates/circuits/poseidon2-air/src/babybear.rs:20-64 (constants = zkhash RC16 rows 0-3,
//               first element of rows 4-16, rows 17-20).
// PARITY UNPINNED items restated from the absent dependencies:
//   * RC16 (HorizenLabs/poseidon2 @ bb476b9, poseidon2_instance_babybear.rs): regenerated here with the
//     published Grain-LFSR procedure (poseidon2_rust_params.sage: field=1, sbox=0, n=31, t=16, R_F=8,
//     R_P=13; (R_F*t + R_P) constants in round order). Self-check: the first row reproduces the 16 words
//     0x69cbb6af 0x46ad93f9 0x60a00f4e 0x6b1297cd ... 0x1ab7202e of the published table
//     (tests/test_constants.py).
//   * internal diagonal d[3..15] (p3-baby-bear BabyBearInternalLayerParameters, INTERNAL_DIAG_MONTY):
//     [-2, 1, 2, 1/2, 3, 4, -1/2, -3, -4, 1/2^8, 1/4, 1/8, 1/2^27, -1/2^8, -1/16, -1/2^27].
#pragma once
#include "babybear.hpp"
#include <cstring>

namespace orc {

struct Poseidon2Constants {
    F ext_init[4][16];
    F partial[13];
    F ext_term[4][16];
    F diag[16];
};

namespace detail {
struct Grain {
    uint8_t s[80];
    int head = 0;
    Grain(unsigned field, unsigned sbox, unsigned n, unsigned t, unsigned rf, unsigned rp) {
        int k = 0;
        auto put = [&](unsigned v, int w) { for (int i = w - 1; i >= 0; i--) s[k++] = (v >> i) & 1; };
        put(field, 2); put(sbox, 4); put(n, 12); put(t, 12); put(rf, 10); put(rp, 10);
        for (int i = 0; i < 30; i++) s[k++] = 1;
        for (int i = 0; i < 160; i++) step();
    }
    inline uint8_t at(int i) const { return s[(head + i) % 80]; }
    uint8_t step() {
        uint8_t nb = at(62) ^ at(51) ^ at(38) ^ at(23) ^ at(13) ^ at(0);
        s[head] = nb;  // pop front, push back
        head = (head + 1) % 80;
        return nb;
    }
    // self-shrinking output: take a pair; if first is 1 emit second, else discard
    uint8_t next_bit() {
        for (;;) {
            uint8_t a = step();
            uint8_t b = step();
            if (a) return b;
        }
    }
    uint32_t next_field(unsigned n, uint32_t p) {
        for (;;) {
            uint32_t v = 0;
            for (unsigned i = 0; i < n; i++) v = (v << 1) | next_bit();
            if (v < p) return v;
        }
    }
};
}  // namespace detail

inline const Poseidon2Constants& poseidon2_constants() {
    static const Poseidon2Constants C = [] {
        Poseidon2Constants c;
        detail::Grain g(1, 0, 31, 16, 8, 13);
        for (int r = 0; r < 4; r++) for (int i = 0; i < 16; i++) c.ext_init[r][i] = F(g.next_field(31, P));
        for (int r = 0; r < 13; r++) c.partial[r] = F(g.next_field(31, P));
        for (int r = 0; r < 4; r++) for (int i = 0; i < 16; i++) c.ext_term[r][i] = F(g.next_field(31, P));
        auto inv2k = [](unsigned k) { return finv(fpow(F(2), k)); };
        F d[16] = {
            -F(2), F(1), F(2), inv2k(1), F(3), F(4), -inv2k(1), -F(3), -F(4),
            inv2k(8), inv2k(2), inv2k(3), inv2k(27), -inv2k(8), -inv2k(4), -inv2k(27)};
        for (int i = 0; i < 16; i++) c.diag[i] = d[i];
        return c;
    }();
    return C;
}

inline F sbox7(F x) { F x2 = x * x, x3 = x2 * x, x4 = x2 * x2; return x3 * x4; }

inline void ext_layer(F s[16]) {
    // 4x4 block M4 = [[2,3,1,1],[1,2,3,1],[1,1,2,3],[3,1,1,2]] on each group of four ...
    for (int b = 0; b < 16; b += 4) {
        F a = s[b], bb = s[b + 1], c = s[b + 2], d = s[b + 3];
        F two(2), three(3);
        s[b] = two * a + three * bb + c + d;
        s[b + 1] = a + two * bb + three * c + d;
        s[b + 2] = a + bb + two * c + three * d;
        s[b + 3] = three * a + bb + c + two * d;
    }
    // ... then add the column sums of the four blocks (circ(2*M4, M4, M4, M4)).
    F sums[4];
    for (int i = 0; i < 16; i++) sums[i & 3] += s[i];
    for (int i = 0; i < 16; i++) s[i] += sums[i & 3];
}

inline void int_layer(F s[16], const F diag[16]) {
    F sum;
    for (int i = 0; i < 16; i++) sum += s[i];
    for (int i = 0; i < 16; i++) s[i] = sum + diag[i] * s[i];
}

inline void poseidon2_permute(F s[16]) {
    const Poseidon2Constants& C = poseidon2_constants();
    ext_layer(s);
    for (int r = 0; r < 4; r++) {
        for (int i = 0; i < 16; i++) s[i] = sbox7(s[i] + C.ext_init[r][i]);
        ext_layer(s);
    }
    for (int r = 0; r < 13; r++) {
        s[0] = sbox7(s[0] + C.partial[r]);
        int_layer(s, C.diag);
    }
    for (int r = 0; r < 4; r++) {
        for (int i = 0; i < 16; i++) s[i] = sbox7(s[i] + C.ext_term[r][i]);
        ext_layer(s);
    }
}

typedef std::array<F, 8> Digest;

// PaddingFreeSponge<Perm,16,8,8> (hints.rs:36; extensions/native/circuit/src/poseidon2/README.md:9-17;
// execution.rs:478-522): overwrite-mode absorb, no padding, permute after every (possibly short) chunk.
struct SpongeBabyBear {
    F st[16];
    int pos = 0;
    void absorb(F x) {
        st[pos++] = x;
        if (pos == 8) { poseidon2_permute(st); pos = 0; }
    }
    Digest finish() {
        if (pos) { poseidon2_permute(st); pos = 0; }
        Digest d;
        for (int i = 0; i < 8; i++) d[i] = st[i];
        return d;
    }
};

// TruncatedPermutation<Perm,2,8,16> (hints.rs:37-38; crates/vm/src/arch/hasher/poseidon2.rs:33-43)
inline Digest compress_babybear(const Digest& l, const Digest& r) {
    F s[16];
    for (int i = 0; i < 8; i++) { s[i] = l[i]; s[8 + i] = r[i]; }
    poseidon2_permute(s);
    Digest d;
    for (int i = 0; i < 8; i++) d[i] = s[i];
    return d;
}

}  // namespace orc

#include "bn254.hpp"  // Sponge / hash_slice / compress of the selected configuration
