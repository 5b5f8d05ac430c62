// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// DuplexChallenger<F, Poseidon2BabyBear<16>, 16, 8> (absent dependency p3-challenger @ 539bbc84; PARITY UNPINNED),
// restated from the in-tree recursive-verifier twin
//   extensions/native/recursion/src/challenger/duplex.rs:27-131
// observe: clears the output buffer, writes the value at sponge_state[input_ptr++], permutes when 8 inputs are buffered;
// sample: permutes first if inputs are pending or no output is left, then pops from the END of state[0..8);
// sample_ext: four samples = coefficients [1,x,x^2,x^3]; sample_bits(n): low n bits of the canonical sample;
// check_witness(bits, w): observe(w); sample_bits(bits) == 0.
#pragma once
#include "poseidon2.hpp"

namespace orc {

struct Challenger {
    F st[16];
    int in_ptr = 0;   // number of pending inputs already written into st[0..in_ptr)
    int out_ptr = 0;  // number of outputs still available: st[0..out_ptr)

    void duplexing() {
        in_ptr = 0;
        poseidon2_permute(st);
        out_ptr = 8;
    }
    void observe(F v) {
        out_ptr = 0;
        st[in_ptr++] = v;
        if (in_ptr == 8) duplexing();
    }
    void observe_digest(const Digest& d) { for (int i = 0; i < 8; i++) observe(d[i]); }
    void observe_ext(const EF& e) { for (int i = 0; i < 4; i++) observe(e.c[i]); }
    F sample() {
        if (in_ptr != 0) duplexing();
        else if (out_ptr == 0) duplexing();
        out_ptr--;
        return st[out_ptr];
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
