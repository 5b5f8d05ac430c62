// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the shipped product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// build, link or call anything under oracle/.
//
// PARITY: UNPINNED at the dependency boundary. The reference implementation of this arithmetic lives
// in Plonky3 (p3-baby-bear, p3-field @ 539bbc84) which is NOT vendored under /root/reference. This
// file restates the published definitions and is anchored on the reference's in-tree call sites:
//   * p = 2^31 - 2^27 + 1, two-adicity 27   (extensions/native/recursion/src/fri/two_adic_pcs.rs:20-22,
//                                            crates/vm/src/arch/vm.rs:608-610)
//   * EF = F[x]/(x^4 - 11), basis [1,x,x^2,x^3]
//       (extensions/native/circuit/src/field_extension/core.rs:24,286-330;
//        crates/circuits/primitives/cuda/include/primitives/constants.h:18-23)
//   * GENERATOR = 31 and two_adic_generator(27) = 31^((p-1)/2^27) = 0x1a427a41 (self-checked in
//     tests/test_constants.py: generator property, exact order, and the 28-entry table recalled from
//     p3-baby-bear is closed under squaring).
//
// Representation: canonical u32 in [0,p) with 64-bit products and '%' — deliberately NOT the
// Montgomery form used by the CUDA product path, so that the two implementations are independent.
#pragma once
#include <cstdint>
#include <cstddef>
#include <vector>
#include <array>

namespace orc {

constexpr uint32_t P = 2013265921u;  // 0x78000001
constexpr uint32_t GENERATOR = 31u;
constexpr uint32_t TWO_ADICITY = 27u;
constexpr uint32_t EXT_W = 11u;  // x^4 = 11

struct F {
    uint32_t v;
    F() : v(0) {}
    explicit constexpr F(uint32_t x) : v(x) {}
    static F from_u64(uint64_t x) { return F((uint32_t)(x % P)); }
    bool operator==(const F& o) const { return v == o.v; }
    bool operator!=(const F& o) const { return v != o.v; }
};
inline F operator+(F a, F b) { uint32_t s = a.v + b.v; return F(s >= P ? s - P : s); }
inline F operator-(F a, F b) { return F(a.v >= b.v ? a.v - b.v : a.v + P - b.v); }
inline F operator-(F a) { return F(a.v ? P - a.v : 0); }
inline F operator*(F a, F b) { return F((uint32_t)(((uint64_t)a.v * b.v) % P)); }
inline F& operator+=(F& a, F b) { a = a + b; return a; }
inline F& operator-=(F& a, F b) { a = a - b; return a; }
inline F& operator*=(F& a, F b) { a = a * b; return a; }
inline F fpow(F b, uint64_t e) {
    F r(1);
    while (e) { if (e & 1) r *= b; b *= b; e >>= 1; }
    return r;
}
inline F finv(F a) { return fpow(a, P - 2); }
// two_adic_generator(bits): generator of the 2^bits subgroup; table root = 31^((p-1)/2^27).
inline F two_adic_generator(unsigned bits) {
    F g = fpow(F(GENERATOR), (uint64_t)(P - 1) >> TWO_ADICITY);
    for (unsigned i = bits; i < TWO_ADICITY; i++) g *= g;
    return g;
}

// Degree-4 binomial extension, x^4 = 11 (field_extension/core.rs:286-330).
struct EF {
    F c[4];
    EF() {}
    explicit EF(F a) { c[0] = a; }
    EF(F a, F b, F cc, F d) { c[0] = a; c[1] = b; c[2] = cc; c[3] = d; }
    bool operator==(const EF& o) const { return c[0] == o.c[0] && c[1] == o.c[1] && c[2] == o.c[2] && c[3] == o.c[3]; }
    bool operator!=(const EF& o) const { return !(*this == o); }
    bool is_zero() const { return !(c[0].v | c[1].v | c[2].v | c[3].v); }
};
inline EF operator+(EF a, const EF& b) { for (int i = 0; i < 4; i++) a.c[i] += b.c[i]; return a; }
inline EF operator-(EF a, const EF& b) { for (int i = 0; i < 4; i++) a.c[i] -= b.c[i]; return a; }
inline EF operator-(EF a) { for (int i = 0; i < 4; i++) a.c[i] = -a.c[i]; return a; }
inline EF operator+(EF a, F b) { a.c[0] += b; return a; }
inline EF operator-(EF a, F b) { a.c[0] -= b; return a; }
inline EF operator*(EF a, F b) { for (int i = 0; i < 4; i++) a.c[i] *= b; return a; }
inline EF operator*(const EF& a, const EF& b) {
    // schoolbook with x^4 -> 11
    F t[7];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) t[i + j] += a.c[i] * b.c[j];
    EF r;
    F w(EXT_W);
    r.c[0] = t[0] + w * t[4];
    r.c[1] = t[1] + w * t[5];
    r.c[2] = t[2] + w * t[6];
    r.c[3] = t[3];
    return r;
}
inline EF& operator+=(EF& a, const EF& b) { a = a + b; return a; }
inline EF& operator-=(EF& a, const EF& b) { a = a - b; return a; }
inline EF& operator*=(EF& a, const EF& b) { a = a * b; return a; }
inline EF epow(EF b, uint64_t e) {
    EF r(F(1));
    while (e) { if (e & 1) r *= b; b *= b; e >>= 1; }
    return r;
}
// Inverse via the norm to the base field: a^-1 = a^(p+p^2+p^3) / a^(1+p+p^2+p^3).
// Frobenius x -> x^p acts on basis x^i as multiplication by 11^(i(p-1)/4).
inline EF frobenius(const EF& a) {
    static const F z = fpow(F(EXT_W), (P - 1) / 4);
    F z2 = z * z, z3 = z2 * z;
    return EF(a.c[0], a.c[1] * z, a.c[2] * z2, a.c[3] * z3);
}
inline EF einv(const EF& a) {
    EF f1 = frobenius(a), f2 = frobenius(f1), f3 = frobenius(f2);
    EF prod = f1 * f2 * f3;
    EF n = a * prod;  // lies in the base field
    F ninv = finv(n.c[0]);
    return prod * ninv;
}
inline EF operator/(const EF& a, const EF& b) { return a * einv(b); }

inline unsigned bitrev(unsigned x, unsigned bits) {
    unsigned r = 0;
    for (unsigned i = 0; i < bits; i++) { r = (r << 1) | ((x >> i) & 1); }
    return r;
}
inline uint64_t bitrev64(uint64_t x, unsigned bits) {
    uint64_t r = 0;
    for (unsigned i = 0; i < bits; i++) { r = (r << 1) | ((x >> i) & 1); }
    return r;
}

}  // namespace orc
