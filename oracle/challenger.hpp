// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// DuplexChallenger<F, Poseidon2BabyBear<16>, 16, 8> (absent dependency p3-challenger @ 539bbc84; PARITY UNPINNED),
// restated from the in-tree recursive-verifier twin
//   extensions/native/recursion/src/challenger/duplex.rs:27-131
// observe: clears the output buffer, writes the value at sponge_state[input_ptr++], permutes when 8 inputs are buffered;
// sample: permutes first if inputs are pending or no output is left, then pops from the END of state[0..8);
// sample_ext: four samples = coefficients [1,x,x^2,x^3]; sample_bits(n): low n bits of the canonical sample;
// check_witness(bits, w): observe(w); sample_bits(bits) == 0.
//
// With the root configuration (hash_kind() == 1) the same interface is MultiField32Challenger<BabyBear, Bn254Fr, Perm, 3, 2>,
// restated from extensions/native/recursion/src/challenger/multi_field32.rs:24-108: observed values are buffered, six of
// them (two elements of three 32-bit-spaced words) trigger a duplexing; a duplexing overwrites the first elements of the
// state with the packed buffer, permutes, and refills the output buffer with the three low 64-bit limbs (mod p) of each of
// the three state elements; sample pops from the END of that buffer; a digest is observed as the three low limbs of its scalar.
#pragma once
#include "poseidon2.hpp"

namespace orc {

struct Challenger {
    F st[16];
    int in_ptr = 0;   // number of pending inputs already written into st[0..in_ptr)
    int out_ptr = 0;  // number of outputs still available: st[0..out_ptr)
    // root configuration
    bool root = hash_kind() != 0;
    Bn bst[3] = {{{0, 0, 0, 0}}, {{0, 0, 0, 0}}, {{0, 0, 0, 0}}};
    F bin[6], bout[9];

    void duplexing() {
        if (root) {
            for (int h = 0; 3 * h < in_ptr; h++) bst[h] = bn::reduce_32(bin + 3 * h, (size_t)std::min(3, in_ptr - 3 * h));
            in_ptr = 0;
            bn::permute(bst);
            for (int k = 0; k < 3; k++) bn::split_32(bst[k], bout + 3 * k);
            out_ptr = 9;
            return;
        }
        in_ptr = 0;
        poseidon2_permute(st);
        out_ptr = 8;
    }
    void observe(F v) {
        out_ptr = 0;
        if (root) {
            bin[in_ptr++] = v;
            if (in_ptr == 6) duplexing();
            return;
        }
        st[in_ptr++] = v;
        if (in_ptr == 8) duplexing();
    }
    void observe_digest(const Digest& d) {
        if (root) {
            for (int j = 0; j < 3; j++) observe(F((uint32_t)((((uint64_t)d[2 * j + 1].v << 32) | d[2 * j].v) % P)));
            return;
        }
        for (int i = 0; i < 8; i++) observe(d[i]);
    }
    void observe_ext(const EF& e) { for (int i = 0; i < 4; i++) observe(e.c[i]); }
    F sample() {
        if (in_ptr != 0) duplexing();
        else if (out_ptr == 0) duplexing();
        out_ptr--;
        return root ? bout[out_ptr] : st[out_ptr];
    }
    EF sample_ext() {
        EF e;
        for (int i = 0; i < 4; i++) e.c[i] = sample();
        return e;
    }
    uint32_t sample_bits(unsigned bits) {
        uint32_t v = sample().v;
        return bits >= 32 ? v : (v & ((1u << bits) - 1));
    }
    bool check_witness(unsigned bits, F w) {
        observe(w);
        return sample_bits(bits) == 0;
    }
    // minimum canonical witness (the reference searches in parallel and returns any valid one)
    F grind(unsigned bits) {
        for (uint32_t w = 0; w < P; w++) {
            Challenger c = *this;
            if (c.check_witness(bits, F(w))) {
                check_witness(bits, F(w));
                return F(w);
            }
        }
        return F(0);
    }
};

}  // namespace orc
