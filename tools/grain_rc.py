"""Poseidon2 round-constant generation (Grain LFSR), as published in the Poseidon/Poseidon2
reference parameter script (HorizenLabs/poseidon2, poseidon2_rust_params.sage).
Used by tools/gen_constants.py; restated independently in C++ in oracle/poseidon2_oracle.cpp.
"""
P = 2013265921  # BabyBear 2^31 - 2^27 + 1

def grain_gen(field, sbox, n, t, rf, rp):
    def bits(v, w):
        return [int(c) for c in bin(v)[2:].zfill(w)]
    s = bits(field, 2) + bits(sbox, 4) + bits(n, 12) + bits(t, 12) + bits(rf, 10) + bits(rp, 10) + [1] * 30
    assert len(s) == 80
    def step():
        nb = s[62] ^ s[51] ^ s[38] ^ s[23] ^ s[13] ^ s[0]
        s.pop(0); s.append(nb)
        return nb
    for _ in range(160):
        step()
    while True:
        nb = step()
        while nb == 0:
            step()
            nb = step()
        yield step()

def round_constants(p=P, n=31, t=16, rf=8, rp=13):
    g = grain_gen(1, 0, n, t, rf, rp)
    def rnd():
        while True:
            v = 0
            for _ in range(n):
                v = (v << 1) | next(g)
            if v < p:
                return v
    num = rf * t + rp
    out = []
    i = 0
    for i in range(num):
        out.append(rnd())
    # layout: rf/2 full rounds (t each), rp partial (1 each), rf/2 full rounds
    ext_init = [out[r * t:(r + 1) * t] for r in range(rf // 2)]
    off = (rf // 2) * t
    partial = out[off:off + rp]
    off += rp
    ext_term = [out[off + r * t: off + (r + 1) * t] for r in range(rf // 2)]
    return ext_init, partial, ext_term

if __name__ == "__main__":
    a, b, c = round_constants()
    print([hex(x) for x in a[0]])
    print([hex(x) for x in b])
    print([hex(x) for x in c[3]])
