// Host side of the Fiat-Shamir transcript: Poseidon2 permutation and DuplexChallenger<F, Perm, 16, 8> on the CPU,
// Montgomery form (shares field.cuh and the generated constants with the device code).
//
// Replaces (reference, external): p3-challenger `DuplexChallenger`, host-side and sequential in the reference as
// well; in-tree twin extensions/native/recursion/src/challenger/duplex.rs:27-131 (observe / sample pop order /
// sample_bits / check_witness). The transcript is a few thousand permutations per proof; the searches that are
// not tiny (proof-of-work grinding) run on the device (prover.cu: pow_grind_kernel).
#pragma once
#include "field.cuh"
#include "poseidon2_constants.cuh"
#include <cstdlib>
#include <array>
#include "bn254.cuh"

// host_poseidon2_avx2.cpp: the same permutation eight lanes at a time
extern "C" int b2_host_has_avx2();
extern "C" void b2_poseidon2_permute_avx2(uint32_t* state16);

namespace b2 {

struct HostPoseidon2 {
    uint32_t ext[8][16] = B2_P2_EXT_ALL;
    uint32_t partial[13] = B2_P2_PARTIAL;
    uint32_t diag[16] = B2_P2_DIAG;

    static Fp sbox(Fp x) { Fp x2 = x * x, x3 = x2 * x, x4 = x2 * x2; return x3 * x4; }
    static void external(Fp* s) {
        for (int b = 0; b < 16; b += 4) {
            Fp x0 = s[b], x1 = s[b + 1], x2 = s[b + 2], x3 = s[b + 3];
            Fp sum = x0 + x1 + x2 + x3;
            s[b] = sum + x0 + x1 + x1;          // 2x0 + 3x1 + x2 + x3
            s[b + 1] = sum + x1 + x2 + x2;      // x0 + 2x1 + 3x2 + x3
            s[b + 2] = sum + x2 + x3 + x3;      // x0 + x1 + 2x2 + 3x3
            s[b + 3] = sum + x3 + x0 + x0;      // 3x0 + x1 + x2 + 2x3
        }
        Fp sums[4];
        for (int i = 0; i < 16; i++) sums[i & 3] += s[i];
        for (int i = 0; i < 16; i++) s[i] += sums[i & 3];
    }
    void permute(Fp* s) const {
        static const bool avx2 = b2_host_has_avx2() != 0 && !getenv("B2S_HOST_SCALAR");
        if (avx2) {
            static_assert(sizeof(Fp) == 4, "Fp is one Montgomery word");
            b2_poseidon2_permute_avx2(reinterpret_cast<uint32_t*>(s));
            return;
        }
        permute_scalar(s);
    }
    void permute_scalar(Fp* s) const {
        external(s);
        for (int r = 0; r < 4; r++) {
            for (int i = 0; i < 16; i++) s[i] = sbox(s[i] + Fp::raw(ext[r][i]));
            external(s);
        }
        for (int r = 0; r < 13; r++) {
            s[0] = sbox(s[0] + Fp::raw(partial[r]));
            Fp sum;
            for (int i = 0; i < 16; i++) sum += s[i];
            for (int i = 0; i < 16; i++) s[i] = sum + s[i] * Fp::raw(diag[i]);
        }
        for (int r = 4; r < 8; r++) {
            for (int i = 0; i < 16; i++) s[i] = sbox(s[i] + Fp::raw(ext[r][i]));
            external(s);
        }
    }
};

// DuplexChallenger<BabyBear, Poseidon2BabyBear<16>, 16, 8>, or -- hash_kind 1, the root configuration --
// MultiField32Challenger<BabyBear, Bn254Fr, Poseidon2Bn254<3>, 3, 2> as the in-tree circuit twin states it
// (extensions/native/recursion/src/challenger/multi_field32.rs:24-108): observed values are buffered, six of them (two
// scalars of three 32-bit-spaced words) trigger a duplexing; a duplexing overwrites the leading scalars with the packed
// buffer, permutes, and refills the outputs with the three low 64-bit limbs (mod p) of each of the three scalars; sample
// pops from the end; a digest is observed as the three low limbs of its scalar.
struct HostChallenger {
    HostPoseidon2 perm;
    Fp st[16];
    int in_ptr = 0, out_ptr = 0;
    uint64_t permutations = 0;
    bool root = false;
    Fr bst[3] = {fr_zero(), fr_zero(), fr_zero()};
    uint32_t bin[6] = {0, 0, 0, 0, 0, 0};   // canonical
    Fp bout[9];

    explicit HostChallenger(int hash_kind = 0) : root(hash_kind != 0) {}

    void duplexing() {
        permutations++;
        if (root) {
            for (int h = 0; 3 * h < in_ptr; h++) {
                uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int k = 0; k < 3 && 3 * h + k < in_ptr; k++) w[k] = bin[3 * h + k];
                bst[h] = bn_reduce_32(w);
            }
            in_ptr = 0;
            poseidon2_bn254_permute(bst);
            for (int k = 0; k < 3; k++) {
                uint32_t f[3];
                bn_split_32(bst[k], f);
                for (int j = 0; j < 3; j++) bout[3 * k + j] = Fp::from_canonical(f[j]);
            }
            out_ptr = 9;
            return;
        }
        in_ptr = 0;
        perm.permute(st);
        out_ptr = 8;
    }
    void observe(Fp v) {
        out_ptr = 0;
        if (root) {
            bin[in_ptr++] = v.to_canonical();
            if (in_ptr == 6) duplexing();
            return;
        }
        st[in_ptr++] = v;
        if (in_ptr == 8) duplexing();
    }
    void observe_canonical(uint32_t c) { observe(Fp::from_canonical(c)); }
    // a digest as it sits in a tree layer: Montgomery words, or (root) the canonical limbs of its scalar
    void observe_digest(const uint32_t* words8) {
        if (root) {
            for (int j = 0; j < 3; j++)
                observe_canonical((uint32_t)((((uint64_t)words8[2 * j + 1] << 32) | words8[2 * j]) % P));
            return;
        }
        for (int i = 0; i < 8; i++) observe(Fp::raw(words8[i]));
    }
    void observe_digest(const std::array<Fp, 8>& d) {
        uint32_t w[8];
        for (int i = 0; i < 8; i++) w[i] = d[i].v;
        observe_digest(w);
    }
    void observe_ext(const Fp4& e) { for (int i = 0; i < 4; i++) observe(e.c[i]); }
    Fp sample() {
        if (in_ptr != 0 || out_ptr == 0) duplexing();
        --out_ptr;
        return root ? bout[out_ptr] : st[out_ptr];
    }
    Fp4 sample_ext() {
        Fp4 e;
        for (int i = 0; i < 4; i++) e.c[i] = sample();
        return e;
    }
    uint32_t sample_bits(unsigned bits) {
        uint32_t v = sample().to_canonical();
        return bits >= 32 ? v : (v & ((1u << bits) - 1));
    }
    bool check_witness(unsigned bits, Fp w) {
        observe(w);
        return sample_bits(bits) == 0;
    }
};

}  // namespace b2
