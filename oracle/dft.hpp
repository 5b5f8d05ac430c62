// ORACLE — TEST INFRASTRUCTURE ONLY (see babybear.hpp header).
//
// Radix-2 DFT over two-adic cosets and the coset low-degree extension performed by
// TwoAdicFriPcs::commit (absent dependency p3-fri/p3-dft @ 539bbc84; PARITY UNPINNED), anchored on:
//   extensions/native/recursion/src/fri/two_adic_pcs.rs:178-186  leaf i of an LDE of height 2^h holds
//        the evaluation at x = GENERATOR * w_h^{bitrev_h(i)}  (bit-reversed row order)
//   extensions/native/recursion/src/fri/domain.rs:142-154        quotient/LDE domain = GENERATOR * H
//   crates/vm/src/system/program/trace.rs:58-63                  pcs.commit(vec![(domain, trace)])
// Definition restated: given evaluations of a polynomial f (deg < N) on shift_in * H_N in natural
// order, produce evaluations of f on GENERATOR * H_{N*2^b}, stored in bit-reversed row order.
#pragma once
#include "babybear.hpp"
#include <vector>

namespace orc {

// in-place iterative Cooley-Tukey: a[k] <- sum_i a[i] * w^(i*k), natural in, natural out.
inline void ntt_inplace(F* a, unsigned logn, F root) {
    size_t n = (size_t)1 << logn;
    for (size_t i = 0; i < n; i++) {
        size_t j = bitrev64(i, logn);
        if (i < j) std::swap(a[i], a[j]);
    }
    for (unsigned s = 1; s <= logn; s++) {
        size_t m = (size_t)1 << s, half = m >> 1;
        F wm = root;
        for (unsigned k = s; k < logn; k++) wm *= wm;  // w^(n/m)
        std::vector<F> tw(half);
        F w(1);
        for (size_t j = 0; j < half; j++) { tw[j] = w; w *= wm; }
        for (size_t k = 0; k < n; k += m)
            for (size_t j = 0; j < half; j++) {
                F t = tw[j] * a[k + j + half];
                F u = a[k + j];
                a[k + j] = u + t;
                a[k + j + half] = u - t;
            }
    }
}

inline void dft(F* a, unsigned logn) { ntt_inplace(a, logn, two_adic_generator(logn)); }
inline void idft(F* a, unsigned logn) {
    size_t n = (size_t)1 << logn;
    ntt_inplace(a, logn, finv(two_adic_generator(logn)));
    F ninv = finv(F((uint32_t)(n % P)));
    for (size_t i = 0; i < n; i++) a[i] *= ninv;
}

// One column: evals on shift_in*H_N (natural order) -> evals on GENERATOR*H_{N<<b}, bit-reversed order.
inline void coset_lde_column(const F* in, unsigned logn, unsigned log_blowup, F shift_in, F* out) {
    size_t n = (size_t)1 << logn;
    unsigned logm = logn + log_blowup;
    size_t m = (size_t)1 << logm;
    std::vector<F> c(m);
    for (size_t i = 0; i < n; i++) c[i] = in[i];
    idft(c.data(), logn);  // coefficients of f'(x) = f(shift_in * x)
    F s = F(GENERATOR) * finv(shift_in);
    F pw(1);
    for (size_t i = 0; i < n; i++) { c[i] *= pw; pw *= s; }
    dft(c.data(), logm);   // natural order evaluations at s * w_m^k
    for (size_t k = 0; k < m; k++) out[bitrev64(k, logm)] = c[k];
}

// Naive O(N^2) evaluation used only by tests to pin the fast transform on tiny sizes.
inline F eval_poly_from_evals_naive(const F* evals, unsigned logn, F shift_in, F x) {
    // Lagrange interpolation over shift_in * H_N
    size_t n = (size_t)1 << logn;
    F g = two_adic_generator(logn);
    F acc;
    F xi = shift_in;
    for (size_t i = 0; i < n; i++) {
        F num(1), den(1);
        F xj = shift_in;
        for (size_t j = 0; j < n; j++) {
            if (j != i) { num *= (x - xj); den *= (xi - xj); }
            xj *= g;
        }
        acc += evals[i] * num * finv(den);
        xi *= g;
    }
    return acc;
}

}  // namespace orc
