"""Independent pure-Python restatement (plain ints) of the small primitives, used to cross-check the C++
oracle on tiny cases. Follows the same in-tree sources as oracle/*.hpp; written separately on purpose."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tools"))
from grain_rc import round_constants  # noqa: E402

P = 2013265921
GEN = 31


def inv(x):
    return pow(x % P, P - 2, P)


def two_adic_generator(bits):
    return pow(pow(GEN, (P - 1) >> 27, P), 1 << (27 - bits), P)


EXT_INIT, PARTIAL, EXT_TERM = round_constants()
DIAG = [(-2) % P, 1, 2, inv(2), 3, 4, (-inv(2)) % P, (-3) % P, (-4) % P, inv(1 << 8), inv(4), inv(8), inv(1 << 27),
        (-inv(1 << 8)) % P, (-inv(16)) % P, (-inv(1 << 27)) % P]
M4 = [[2, 3, 1, 1], [1, 2, 3, 1], [1, 1, 2, 3], [3, 1, 1, 2]]


def ext_layer(s):
    # circ(2*M4, M4, M4, M4) written as a dense 16x16 matrix product (linearlayer.cuh:13-55)
    out = [0] * 16
    for bi in range(4):
        for i in range(4):
            acc = 0
            for bj in range(4):
                mult = 2 if bi == bj else 1
                for j in range(4):
                    acc += mult * M4[i][j] * s[4 * bj + j]
            out[4 * bi + i] = acc % P
    return out


def int_layer(s):
    tot = sum(s) % P
    return [(tot + DIAG[i] * s[i]) % P for i in range(16)]


def permute(s):
    s = ext_layer(list(s))
    for r in range(4):
        s = ext_layer([pow((s[i] + EXT_INIT[r][i]) % P, 7, P) for i in range(16)])
    for r in range(13):
        s[0] = pow((s[0] + PARTIAL[r]) % P, 7, P)
        s = int_layer(s)
    for r in range(4):
        s = ext_layer([pow((s[i] + EXT_TERM[r][i]) % P, 7, P) for i in range(16)])
    return s


def hash_slice(xs):
    st = [0] * 16
    xs = list(xs)
    for i in range(0, len(xs), 8):
        chunk = xs[i:i + 8]
        st[:len(chunk)] = chunk
        st = permute(st)
    return st[:8]


def compress(l, r):
    return permute(list(l) + list(r))[:8]


def mmcs_root(mats):
    """mats: list of list-of-rows (height power of two). README.md:9-17 procedure."""
    order = sorted(range(len(mats)), key=lambda i: -len(mats[i]))
    hmax = len(mats[order[0]])

    def rowhash(h, r):
        row = []
        for i in order:
            if len(mats[i]) == h:
                row += list(mats[i][r])
        return hash_slice(row)

    heights = {len(m) for m in mats}
    layer = [rowhash(hmax, r) for r in range(hmax)]
    h = hmax
    while h > 1:
        h //= 2
        nxt = [compress(layer[2 * i], layer[2 * i + 1]) for i in range(h)]
        if h in heights:
            nxt = [compress(nxt[i], rowhash(h, i)) for i in range(h)]
        layer = nxt
    return layer[0]


def coset_lde_naive(col, log_blowup, shift_in=1):
    """O(N^2): interpolate evals over shift_in*H_N, evaluate on GEN*H_{N<<b}, bit-reversed order."""
    n = len(col)
    logn = n.bit_length() - 1
    g = two_adic_generator(logn)
    # coefficients by inverse DFT definition
    ninv = inv(n)
    ginv = inv(g)
    sinv = inv(shift_in)
    coef = []
    for k in range(n):
        acc = 0
        for i in range(n):
            acc += col[i] * pow(ginv, i * k, P)
        coef.append(acc * ninv % P * pow(sinv, k, P) % P)
    m = n << log_blowup
    logm = logm_ = logn + log_blowup
    w = two_adic_generator(logm)
    out = [0] * m
    for k in range(m):
        x = GEN * pow(w, k, P) % P
        acc = 0
        for c in reversed(coef):
            acc = (acc * x + c) % P
        br = int(format(k, "0%db" % logm_)[::-1], 2) if logm_ else 0
        out[br] = acc
    return out


# ---- root configuration: Poseidon2 over the BN254 scalar field, width 3 (oracle/bn254.hpp names the in-tree sources) ----
BN_R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
BN_EXT_INIT, BN_PARTIAL, BN_EXT_TERM = round_constants(BN_R, 254, 3, 8, 56)


def bn_permute(s):
    def external(v):
        t = sum(v) % BN_R
        return [(x + t) % BN_R for x in v]

    s = external(list(s))
    for rc in BN_EXT_INIT:
        s = external([pow((x + k) % BN_R, 5, BN_R) for x, k in zip(s, rc)])
    for k in BN_PARTIAL:
        s[0] = pow((s[0] + k) % BN_R, 5, BN_R)
        t = sum(s) % BN_R
        s = [(s[0] + t) % BN_R, (s[1] + t) % BN_R, (2 * s[2] + t) % BN_R]
    for rc in BN_EXT_TERM:
        s = external([pow((x + k) % BN_R, 5, BN_R) for x, k in zip(s, rc)])
    return s


def bn_reduce_32(vals):
    return sum(v << (32 * i) for i, v in enumerate(vals)) % BN_R


def bn_split_32(x, n):
    return [((x >> (64 * j)) & 0xFFFFFFFFFFFFFFFF) % P for j in range(n)]


def bn_hash_slice(xs):
    """outer_poseidon2.rs:26-50: blocks of 16 values, 8 per element, overwrite mode, no padding."""
    xs = list(xs)
    st = [0, 0, 0]
    for b in range(0, len(xs), 16):
        block = xs[b:b + 16]
        for h in range(0, len(block), 8):
            st[h // 8] = bn_reduce_32(block[h:h + 8])
        st = bn_permute(st)
    return st[0]


def bn_compress(l, r):
    return bn_permute([l, r, 0])[0]


def bn_limbs(x):
    return [(x >> (32 * i)) & 0xFFFFFFFF for i in range(8)]


def bn_from_limbs(w):
    return sum(int(v) << (32 * i) for i, v in enumerate(w))


class BnChallenger:
    """challenger/multi_field32.rs:24-108."""

    def __init__(self):
        self.state = [0, 0, 0]
        self.inp = []
        self.out = []

    def duplexing(self):
        assert len(self.inp) <= 6
        for i in range(0, len(self.inp), 3):
            self.state[i // 3] = bn_reduce_32(self.inp[i:i + 3])
        self.inp = []
        self.state = bn_permute(self.state)
        self.out = []
        for x in self.state:
            self.out += bn_split_32(x, 3)

    def observe(self, v):
        self.out = []
        self.inp.append(v)
        if len(self.inp) == 6:
            self.duplexing()

    def observe_digest(self, x):
        for f in bn_split_32(x, 3):
            self.observe(f)

    def sample(self):
        if self.inp or not self.out:
            self.duplexing()
        return self.out.pop()
