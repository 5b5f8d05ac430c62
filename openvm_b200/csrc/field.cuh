// BabyBear (p = 2^31 - 2^27 + 1) and its degree-4 extension F[x]/(x^4 - 11) on 32-bit lanes.
//
// Replaces (reference, external): p3-baby-bear / p3-monty-31 `MontyField31<BabyBearParameters>` and
// p3-field `BinomialExtensionField<BabyBear,4>`; in-tree anchors:
//   extensions/native/recursion/src/fri/two_adic_pcs.rs:20-22 (two-adicity 27),
//   extensions/native/circuit/src/field_extension/core.rs:24,286-330 (x^4 = 11, basis [1,x,x^2,x^3]).
//
// Device values are kept in Montgomery form (R = 2^32) everywhere in HBM, which is also the in-memory
// form of Plonky3's `BabyBear`, so a Rust host can hand `RowMajorMatrix<BabyBear>` memory over
// unchanged (repr = B2S_REPR_MONTY). Canonical u32 is accepted/produced at the ABI on request.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define B2_HD __host__ __device__ __forceinline__
#define B2_D __device__ __forceinline__
#else
#define B2_HD inline
#define B2_D inline
#endif

namespace b2 {

constexpr uint32_t P = 0x78000001u;          // 2013265921
constexpr uint32_t P_INV = 0x88000001u;      // p^-1 mod 2^32   (p = 1 + 15*2^27, (1+a)^-1 = 1-a since a^2 = 0 mod 2^32)
constexpr uint32_t P_NEG_INV = 0x77FFFFFFu;  // -p^-1 mod 2^32
constexpr uint32_t R_MOD_P = 0x0FFFFFFEu;    // 2^32 mod p  (Montgomery form of 1)
constexpr uint32_t R2_MOD_P = 0x45DDDDE3u;   // 2^64 mod p  (to_monty multiplier), checked by static_assert below
constexpr uint32_t GENERATOR = 31u;
constexpr unsigned TWO_ADICITY = 27;

// ---- compile-time helpers (host side constant derivation; canonical arithmetic) ------------------------
constexpr uint32_t c_mul(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) % P); }
constexpr uint32_t c_pow(uint32_t b, uint64_t e) {
    uint32_t r = 1;
    while (e) { if (e & 1) r = c_mul(r, b); b = c_mul(b, b); e >>= 1; }
    return r;
}
constexpr uint32_t c_inv(uint32_t a) { return c_pow(a, P - 2); }
constexpr uint32_t c_to_monty(uint32_t a) { return (uint32_t)((((uint64_t)a) << 32) % P); }
static_assert(c_to_monty(1) == R_MOD_P, "R mod p");
static_assert(c_to_monty(R_MOD_P) == R2_MOD_P, "R^2 mod p");
static_assert((uint32_t)(P * P_INV) == 1u, "p^-1");

// ---- Montgomery arithmetic ------------------------------------------------------------------------------
// reduce a 64-bit t < 2^32 * p to t * 2^-32 mod p, result in [0, p)
B2_HD uint32_t monty_reduce(uint64_t t) {
    uint32_t m = (uint32_t)t * P_NEG_INV;
    uint64_t u = t + (uint64_t)m * P;  // low word is 0; t < 2^32 p  =>  u < 2^33 p < 2^64
    uint32_t r = (uint32_t)(u >> 32);  // in [0, 2p)
    uint32_t s = r - P;
    return s < r ? s : r;              // unsigned min trick: r < p => s wraps above r
}
// as above but leaves the result in [0, 2p)
B2_HD uint32_t monty_reduce_lazy(uint64_t t) {
    uint32_t m = (uint32_t)t * P_NEG_INV;
    uint64_t u = t + (uint64_t)m * P;
    return (uint32_t)(u >> 32);
}

struct Fp {
    uint32_t v;  // Montgomery form, canonical range [0,p)
    B2_HD Fp() : v(0) {}
    B2_HD explicit constexpr Fp(uint32_t raw, int) : v(raw) {}
    static B2_HD Fp raw(uint32_t m) { return Fp(m, 0); }
    static B2_HD Fp zero() { return Fp(0u, 0); }
    static B2_HD Fp one() { return Fp(R_MOD_P, 0); }
    static B2_HD Fp from_canonical(uint32_t c) { return Fp(monty_reduce((uint64_t)c * R2_MOD_P), 0); }
    B2_HD uint32_t to_canonical() const { return monty_reduce((uint64_t)v); }
    B2_HD bool operator==(Fp o) const { return v == o.v; }
    B2_HD bool operator!=(Fp o) const { return v != o.v; }
};

B2_HD Fp operator+(Fp a, Fp b) {
    uint32_t s = a.v + b.v;  // < 2p < 2^32
    uint32_t t = s - P;
    return Fp::raw(t < s ? t : s);
}
B2_HD Fp operator-(Fp a, Fp b) {
    uint32_t d = a.v - b.v;
    uint32_t t = d + P;
    return Fp::raw(t < d ? t : d);
}
B2_HD Fp operator-(Fp a) { return Fp::raw(a.v ? P - a.v : 0u); }
B2_HD Fp operator*(Fp a, Fp b) { return Fp::raw(monty_reduce((uint64_t)a.v * b.v)); }
B2_HD Fp& operator+=(Fp& a, Fp b) { a = a + b; return a; }
B2_HD Fp& operator-=(Fp& a, Fp b) { a = a - b; return a; }
B2_HD Fp& operator*=(Fp& a, Fp b) { a = a * b; return a; }
B2_HD Fp fp_double(Fp a) { return a + a; }
B2_HD Fp fp_pow(Fp b, uint64_t e) {
    Fp r = Fp::one();
    while (e) { if (e & 1) r *= b; b *= b; e >>= 1; }
    return r;
}
B2_HD Fp fp_inv(Fp a) { return fp_pow(a, P - 2); }

// Montgomery constants derived at compile time
constexpr uint32_t M_GENERATOR = c_to_monty(GENERATOR);
constexpr uint32_t M_EXT_W = c_to_monty(11);
constexpr uint32_t C_ROOT_27 = c_pow(GENERATOR, (uint64_t)(P - 1) >> TWO_ADICITY);  // 0x1a427a41
static_assert(C_ROOT_27 == 0x1a427a41u, "two_adic_generator(27)");

// two_adic_generator(bits) in Montgomery form (host or device; log-time)
B2_HD Fp two_adic_generator(unsigned bits) {
    Fp g = Fp::raw(c_to_monty(C_ROOT_27));
    for (unsigned i = bits; i < TWO_ADICITY; i++) g *= g;
    return g;
}

// ---- degree-4 extension --------------------------------------------------------------------------------
struct Fp4 {
    Fp c[4];
    B2_HD Fp4() {}
    B2_HD explicit Fp4(Fp a) { c[0] = a; }
    B2_HD Fp4(Fp a, Fp b, Fp cc, Fp d) { c[0] = a; c[1] = b; c[2] = cc; c[3] = d; }
    static B2_HD Fp4 zero() { return Fp4(); }
    static B2_HD Fp4 one() { return Fp4(Fp::one()); }
    B2_HD bool operator==(const Fp4& o) const {
        return c[0] == o.c[0] && c[1] == o.c[1] && c[2] == o.c[2] && c[3] == o.c[3];
    }
    B2_HD bool operator!=(const Fp4& o) const { return !(*this == o); }
};
B2_HD Fp4 operator+(Fp4 a, const Fp4& b) { for (int i = 0; i < 4; i++) a.c[i] += b.c[i]; return a; }
B2_HD Fp4 operator-(Fp4 a, const Fp4& b) { for (int i = 0; i < 4; i++) a.c[i] -= b.c[i]; return a; }
B2_HD Fp4 operator-(Fp4 a) { for (int i = 0; i < 4; i++) a.c[i] = -a.c[i]; return a; }
B2_HD Fp4 operator+(Fp4 a, Fp b) { a.c[0] += b; return a; }
B2_HD Fp4 operator-(Fp4 a, Fp b) { a.c[0] -= b; return a; }
B2_HD Fp4 operator*(Fp4 a, Fp b) { for (int i = 0; i < 4; i++) a.c[i] *= b; return a; }
// schoolbook product with x^4 -> 11 (16 base multiplications + 3 by the constant 11)
B2_HD Fp4 operator*(const Fp4& a, const Fp4& b) {
    const uint32_t a0 = a.c[0].v, a1 = a.c[1].v, a2 = a.c[2].v, a3 = a.c[3].v;
    const uint32_t b0 = b.c[0].v, b1 = b.c[1].v, b2 = b.c[2].v, b3 = b.c[3].v;
    Fp t4 = Fp::raw(monty_reduce((uint64_t)a1 * b3)) + Fp::raw(monty_reduce((uint64_t)a2 * b2)) +
            Fp::raw(monty_reduce((uint64_t)a3 * b1));
    Fp t5 = Fp::raw(monty_reduce((uint64_t)a2 * b3)) + Fp::raw(monty_reduce((uint64_t)a3 * b2));
    Fp t6 = Fp::raw(monty_reduce((uint64_t)a3 * b3));
    Fp w = Fp::raw(M_EXT_W);
    Fp4 r;
    r.c[0] = Fp::raw(monty_reduce((uint64_t)a0 * b0)) + w * t4;
    r.c[1] = Fp::raw(monty_reduce((uint64_t)a0 * b1)) + Fp::raw(monty_reduce((uint64_t)a1 * b0)) + w * t5;
    r.c[2] = Fp::raw(monty_reduce((uint64_t)a0 * b2)) + Fp::raw(monty_reduce((uint64_t)a1 * b1)) +
             Fp::raw(monty_reduce((uint64_t)a2 * b0)) + w * t6;
    r.c[3] = Fp::raw(monty_reduce((uint64_t)a0 * b3)) + Fp::raw(monty_reduce((uint64_t)a1 * b2)) +
             Fp::raw(monty_reduce((uint64_t)a2 * b1)) + Fp::raw(monty_reduce((uint64_t)a3 * b0));
    return r;
}
B2_HD Fp4& operator+=(Fp4& a, const Fp4& b) { a = a + b; return a; }
B2_HD Fp4& operator-=(Fp4& a, const Fp4& b) { a = a - b; return a; }
B2_HD Fp4& operator*=(Fp4& a, const Fp4& b) { a = a * b; return a; }
B2_HD Fp4 fp4_pow(Fp4 b, uint64_t e) {
    Fp4 r = Fp4::one();
    while (e) { if (e & 1) r *= b; b *= b; e >>= 1; }
    return r;
}
// Frobenius x -> x^p: basis x^i picks up 11^(i(p-1)/4)
constexpr uint32_t M_FROB1 = c_to_monty(c_pow(11, (P - 1) / 4));
constexpr uint32_t M_FROB2 = c_to_monty(c_mul(c_pow(11, (P - 1) / 4), c_pow(11, (P - 1) / 4)));
constexpr uint32_t M_FROB3 = c_to_monty(c_pow(c_pow(11, (P - 1) / 4), 3));
B2_HD Fp4 fp4_frobenius(const Fp4& a) {
    return Fp4(a.c[0], a.c[1] * Fp::raw(M_FROB1), a.c[2] * Fp::raw(M_FROB2), a.c[3] * Fp::raw(M_FROB3));
}
// inverse through the norm to the base field
B2_HD Fp4 fp4_inv(const Fp4& a) {
    Fp4 f1 = fp4_frobenius(a), f2 = fp4_frobenius(f1), f3 = fp4_frobenius(f2);
    Fp4 prod = f1 * f2 * f3;
    Fp4 n = a * prod;
    return prod * fp_inv(n.c[0]);
}

// two inverses for one base-field inversion (the exponentiation is most of fp4_inv)
B2_HD void fp4_inv_pair(const Fp4& a, const Fp4& b, Fp4& ia, Fp4& ib) {
    Fp4 a1 = fp4_frobenius(a), a2 = fp4_frobenius(a1), a3 = fp4_frobenius(a2);
    Fp4 b1 = fp4_frobenius(b), b2 = fp4_frobenius(b1), b3 = fp4_frobenius(b2);
    const Fp4 pa = a1 * a2 * a3, pb = b1 * b2 * b3;
    const Fp na = (a * pa).c[0], nb = (b * pb).c[0];
    const Fp t = fp_inv(na * nb);
    ia = pa * (nb * t);
    ib = pb * (na * t);
}

B2_HD uint32_t bitrev32(uint32_t x, unsigned bits) {
#if defined(__CUDA_ARCH__)
    return bits ? (__brev(x) >> (32 - bits)) : 0u;
#else
    uint32_t r = 0;
    for (unsigned i = 0; i < bits; i++) r = (r << 1) | ((x >> i) & 1);
    return r;
#endif
}

}  // namespace b2
