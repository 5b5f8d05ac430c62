// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// The hash side of BabyBearPoseidon2RootConfig (crates/continuations/src/lib.rs:17; crates/sdk/src/prover/root.rs):
// Poseidon2 over the BN254 scalar field with width 3, the multi-field sponge / compression / challenger built on it.
// The Rust types live in absent dependencies (openvm-stark-sdk config/baby_bear_poseidon2_root.rs, p3-bn254-fr,
// p3-symmetric MultiField32PaddingFreeSponge, p3-challenger MultiField32Challenger @ Plonky3 539bbc84); restated from the
// in-tree twins that the halo2 verifier is built from:
//   permutation     extensions/native/compiler/src/constraints/halo2/poseidon2_perm.rs:52-215
//                   (initial circ(2,1,1); 4 full rounds; 56 partial rounds with J + diag(1,1,2); 4 full rounds; x^5)
//   parameters      extensions/native/compiler/src/constraints/halo2/compiler.rs:39-85 (zkhash RC3, MAT_DIAG3_M_1; the
//                   partial rounds use lane 0 of rows 4..59)
//   p2_hash         extensions/native/recursion/src/outer_poseidon2.rs:26-50  (8 = 254/31 words per element, rate 2)
//   p2_compress     extensions/native/recursion/src/outer_poseidon2.rs:52-63
//   reduce_32       extensions/native/recursion/src/utils.rs:42-51
//   split_32        extensions/native/recursion/src/utils.rs:53-58 + compiler.rs:603-632 (64-bit limbs, each mod p)
//   challenger      extensions/native/recursion/src/challenger/multi_field32.rs:24-108 (3 = 254/64 words per element)
// Pinning: RC3 is regenerated with the published Grain LFSR (field=1, sbox=0, n=254, t=3, R_F=8, R_P=56) and the
// permutation reproduces the published known answer of the Poseidon2 reference implementation for input (0, 1, 2)
// (tests/test_bn254.py). The sponge and the challenger have no golden vector in the tree: PARITY UNPINNED.
//
// Arithmetic: four 64-bit limbs, Montgomery form, products through unsigned __int128 (the product code uses eight 32-bit
// limbs; the two share nothing).
#pragma once
#include "poseidon2.hpp"   // included from the end of poseidon2.hpp: Digest, detail::Grain, SpongeBabyBear are defined by now
#include <algorithm>

namespace orc {

struct Bn {
    uint64_t l[4];
    bool operator==(const Bn& o) const { return l[0] == o.l[0] && l[1] == o.l[1] && l[2] == o.l[2] && l[3] == o.l[3]; }
};

namespace bn {
typedef unsigned __int128 u128;
static const uint64_t MOD[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull, 0x30644e72e131a029ull};

inline bool geq_mod(const uint64_t t[4]) {
    for (int i = 3; i >= 0; i--) {
        if (t[i] > MOD[i]) return true;
        if (t[i] < MOD[i]) return false;
    }
    return true;
}
inline void sub_mod(uint64_t t[4]) {
    u128 borrow = 0;
    for (int i = 0; i < 4; i++) {
        u128 x = (u128)t[i] - MOD[i] - borrow;
        t[i] = (uint64_t)x;
        borrow = (x >> 64) & 1;
    }
}
inline Bn add(const Bn& a, const Bn& b) {
    Bn r;
    u128 c = 0;
    for (int i = 0; i < 4; i++) {
        c += (u128)a.l[i] + b.l[i];
        r.l[i] = (uint64_t)c;
        c >>= 64;
    }
    if (geq_mod(r.l)) sub_mod(r.l);
    return r;
}
inline uint64_t neg_inv64() {  // -MOD^-1 mod 2^64 by Newton iteration
    uint64_t x = 1;
    for (int i = 0; i < 6; i++) x *= 2 - MOD[0] * x;
    return (uint64_t)0 - x;
}
// Montgomery product: full 512-bit product, then four reduction steps (a b < MOD 2^256)
inline Bn mul(const Bn& a, const Bn& b) {
    static const uint64_t NINV = neg_inv64();
    uint64_t t[9] = {0};
    for (int i = 0; i < 4; i++) {
        u128 c = 0;
        for (int j = 0; j < 4; j++) {
            c += (u128)a.l[i] * b.l[j] + t[i + j];
            t[i + j] = (uint64_t)c;
            c >>= 64;
        }
        t[i + 4] = (uint64_t)c;
    }
    for (int i = 0; i < 4; i++) {
        const uint64_t m = t[i] * NINV;
        u128 c = 0;
        for (int j = 0; j < 4; j++) {
            c += (u128)m * MOD[j] + t[i + j];
            t[i + j] = (uint64_t)c;
            c >>= 64;
        }
        for (int k = i + 4; k < 9 && c; k++) {
            c += t[k];
            t[k] = (uint64_t)c;
            c >>= 64;
        }
    }
    Bn r = {{t[4], t[5], t[6], t[7]}};
    if (t[8] || geq_mod(r.l)) sub_mod(r.l);
    return r;
}
inline const Bn& r2() {  // 2^512 mod MOD: double 1 five hundred and twelve times
    static const Bn v = [] {
        Bn x = {{1, 0, 0, 0}};
        for (int i = 0; i < 512; i++) x = add(x, x);
        return x;
    }();
    return v;
}
// canonical 256-bit integer (any value below 2^256) -> element
inline Bn from_words(const uint64_t w[4]) {
    Bn a = {{w[0], w[1], w[2], w[3]}};
    return mul(a, r2());
}
inline void to_words(const Bn& a, uint64_t w[4]) {
    Bn one = {{1, 0, 0, 0}};
    Bn c = mul(a, one);
    for (int i = 0; i < 4; i++) w[i] = c.l[i];
}

struct Constants {
    Bn ext[8][3];
    Bn partial[56];
};
inline const Constants& constants() {
    static const Constants C = [] {
        Constants c;
        detail::Grain g(1, 0, 254, 3, 8, 56);
        auto next = [&]() {
            for (;;) {
                uint64_t w[4] = {0, 0, 0, 0};
                for (int i = 0; i < 254; i++) {  // most significant bit first
                    w[3] = (w[3] << 1) | (w[2] >> 63);
                    w[2] = (w[2] << 1) | (w[1] >> 63);
                    w[1] = (w[1] << 1) | (w[0] >> 63);
                    w[0] = (w[0] << 1) | g.next_bit();
                }
                if (!geq_mod(w)) return from_words(w);
            }
        };
        for (int r = 0; r < 4; r++) for (int i = 0; i < 3; i++) c.ext[r][i] = next();
        for (int r = 0; r < 56; r++) c.partial[r] = next();
        for (int r = 4; r < 8; r++) for (int i = 0; i < 3; i++) c.ext[r][i] = next();
        return c;
    }();
    return C;
}
inline Bn pow5(const Bn& x) {
    Bn x2 = mul(x, x), x4 = mul(x2, x2);
    return mul(x, x4);
}
inline void permute(Bn s[3]) {
    const Constants& C = constants();
    auto external = [&]() {
        Bn t = add(add(s[0], s[1]), s[2]);
        for (int i = 0; i < 3; i++) s[i] = add(s[i], t);
    };
    external();
    for (int r = 0; r < 4; r++) {
        for (int i = 0; i < 3; i++) s[i] = pow5(add(s[i], C.ext[r][i]));
        external();
    }
    for (int r = 0; r < 56; r++) {
        s[0] = pow5(add(s[0], C.partial[r]));
        Bn t = add(add(s[0], s[1]), s[2]);
        s[0] = add(s[0], t);
        s[1] = add(s[1], t);
        s[2] = add(add(s[2], s[2]), t);
    }
    for (int r = 4; r < 8; r++) {
        for (int i = 0; i < 3; i++) s[i] = pow5(add(s[i], C.ext[r][i]));
        external();
    }
}
// sum_k v_k 2^(32 k) for up to eight canonical BabyBear values
inline Bn reduce_32(const F* v, size_t n) {
    uint64_t w[4] = {0, 0, 0, 0};
    for (size_t k = 0; k < n; k++) w[k / 2] |= (uint64_t)v[k].v << (32 * (k & 1));
    return from_words(w);
}
inline void split_32(const Bn& x, F out[3]) {
    uint64_t w[4];
    to_words(x, w);
    for (int j = 0; j < 3; j++) out[j] = F((uint32_t)(w[j] % P));
}
// a digest is one element; it travels as the eight 32-bit limbs of its canonical value in a Digest
inline Digest to_digest(const Bn& x) {
    uint64_t w[4];
    to_words(x, w);
    Digest d;
    for (int i = 0; i < 8; i++) d[i].v = (uint32_t)(w[i / 2] >> (32 * (i & 1)));
    return d;
}
inline Bn from_digest(const Digest& d) {
    uint64_t w[4];
    for (int i = 0; i < 4; i++) w[i] = (uint64_t)d[2 * i].v | ((uint64_t)d[2 * i + 1].v << 32);
    return from_words(w);
}
}  // namespace bn

// MultiField32PaddingFreeSponge<BabyBear, Bn254Fr, Perm, 3, 16, 1> fed one value at a time
struct SpongeBn {
    Bn st[3];
    F buf[16];
    int pos = 0;
    SpongeBn() { for (auto& s : st) s = Bn{{0, 0, 0, 0}}; }
    void flush() {
        for (int h = 0; 8 * h < pos; h++) st[h] = bn::reduce_32(buf + 8 * h, (size_t)std::min(8, pos - 8 * h));
        bn::permute(st);
        pos = 0;
    }
    void absorb(F x) {
        buf[pos++] = x;
        if (pos == 16) flush();
    }
    Digest finish() {
        if (pos) flush();
        return bn::to_digest(st[0]);
    }
};
inline Digest compress_bn(const Digest& l, const Digest& r) {
    Bn s[3] = {bn::from_digest(l), bn::from_digest(r), Bn{{0, 0, 0, 0}}};
    bn::permute(s);
    return bn::to_digest(s[0]);
}


// ---- the configuration in force (oracle_capi: orc_set_hash) ----------------------------------------------------------
// 0: BabyBearPoseidon2Config (digests of 8 BabyBear words), 1: BabyBearPoseidon2RootConfig (one BN254 scalar)
inline int& hash_kind() {
    static int kind = 0;
    return kind;
}
struct Sponge {
    SpongeBabyBear a;
    SpongeBn b;
    bool root = hash_kind() != 0;
    void absorb(F x) { root ? b.absorb(x) : a.absorb(x); }
    Digest finish() { return root ? b.finish() : a.finish(); }
};
inline Digest hash_slice(const F* x, size_t n) {
    Sponge s;
    for (size_t i = 0; i < n; i++) s.absorb(x[i]);
    return s.finish();
}
inline Digest compress(const Digest& l, const Digest& r) { return hash_kind() ? compress_bn(l, r) : compress_babybear(l, r); }

}  // namespace orc
