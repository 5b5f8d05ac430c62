"""Proof -> hint stream of the recursive verifier (second half).

Three independent pieces must agree: the product's b2s_proof_hint_stream (C ABI, host-only), the oracle's restatement of
`Hintable::write` (oracle/hints.hpp), and the READER below, which consumes a stream in the order the reference's
`Hintable::read` implementations consume it (extensions/native/recursion/src/hints.rs:272-444, fri/hints.rs:22-163) and
re-encodes what it read into the proof wire format (crates/sdk/src/codec.rs): stream -> bytes must give back the proof.
Known answers of the reference's own tests (hints.rs:455-531) are checked on the vector forms they pin."""
import ctypes as C
import struct

import numpy as np
import pytest

from openvm_b200 import airs
from oracle_binding import unflatten_stream
from test_oracle_stark import build_key, interaction_traces

P = 2013265921


def product_stream(proof):
    from openvm_b200 import _lib
    lib = _lib.load()
    words, n = C.POINTER(C.c_uint32)(), C.c_size_t()
    rc = lib.b2s_proof_hint_stream(None, proof, len(proof), C.byref(words), C.byref(n))
    if rc != 0:
        return rc
    flat = np.ctypeslib.as_array(words, shape=(n.value,)).copy()
    lib.b2s_free(words)
    return unflatten_stream(flat)


class Reader:
    """hint_var / hint_felt / hint_ext / hint_vars / hint_felts / hint_exts of the reference's builder, over a stream"""

    def __init__(self, stream):
        self.s, self.i = stream, 0

    def vec(self):
        v = self.s[self.i]
        self.i += 1
        return v

    def var(self):
        v = self.vec()
        assert len(v) == 1
        return v[0]

    def ext(self):
        return [self.var() for _ in range(4)]

    def exts(self):              # Vec<Challenge>: [len], flat
        n = self.var()
        flat = self.vec()
        assert len(flat) == 4 * n
        return [flat[4 * k:4 * k + 4] for k in range(n)]

    def digest(self):
        return [self.var() for _ in range(8)]

    def digests(self):
        return [self.digest() for _ in range(self.var())]

    def hint_slice(self, unit):  # read_hint_slice: length, then the words
        n = self.var()
        flat = self.vec()
        assert len(flat) == unit * n
        return flat

    def adj(self):
        return self.exts(), self.exts()

    def adjs(self):
        return [self.adj() for _ in range(self.var())]


def stream_to_wire(stream, input_row_widths):
    """Reads a Proof the way StarkProofVariable is read and writes codec.rs bytes. `input_row_widths[q][b]`: widths of the
    rows of batch opening b (the hint stream flattens them; the wire keeps them apart)."""
    r = Reader(stream)
    main, after, quotient = r.digests(), r.digests(), r.digest()
    fri_commits = r.digests()
    queries = []
    for qi in range(r.var()):
        inputs = []
        for bi in range(r.var()):
            flat = r.hint_slice(1)
            inputs.append((flat, r.hint_slice(8)))
        steps = []
        for _ in range(r.var()):
            sib = r.ext()
            steps.append((sib, r.hint_slice(8)))
        queries.append((inputs, steps))
    final_poly = r.exts()
    pow_witness = r.var()
    prep = r.adjs()
    opened_main = [r.adjs() for _ in range(r.var())]
    opened_quot = [[r.exts() for _ in range(r.var())] for _ in range(r.var())]
    opened_after = [r.adjs() for _ in range(r.var())]
    per_air = []
    for _ in range(r.var()):
        air_id, log_degree = r.var(), r.var()
        exposed = [r.exts() for _ in range(r.var())]
        per_air.append((air_id, log_degree, exposed, r.vec()))
    perm = r.vec()
    logup = r.var()
    assert r.i == len(stream)

    out = bytearray()
    u32 = lambda x: out.extend(struct.pack("<I", x))
    words = lambda ws: [u32(w) for w in ws]
    def efs(v):
        u32(len(v)); [words(e) for e in v]
    def digs(v):
        u32(len(v)); [words(d) for d in v]
    def adjs(v):
        u32(len(v))
        for loc, nxt in v:
            efs(loc); efs(nxt)
    u32(1)
    digs(main); digs(after); words(quotient)
    digs(fri_commits)
    u32(len(queries))
    for qi, (inputs, steps) in enumerate(queries):
        u32(len(inputs))
        for bi, (flat, path) in enumerate(inputs):
            widths = input_row_widths[qi][bi]
            assert sum(widths) == len(flat)
            u32(len(widths))
            off = 0
            for w in widths:
                u32(w); words(flat[off:off + w]); off += w
            u32(len(path) // 8); words(path)
        u32(len(steps))
        for sib, path in steps:
            words(sib); u32(len(path) // 8); words(path)
    efs(final_poly); u32(pow_witness)
    adjs(prep)
    u32(len(opened_main)); [adjs(p) for p in opened_main]
    u32(len(opened_after)); [adjs(p) for p in opened_after]
    u32(len(opened_quot))
    for air in opened_quot:
        u32(len(air)); [efs(c) for c in air]
    u32(len(per_air))
    for air_id, lg, exposed, pvs in per_air:
        u32(air_id); u32(1 << lg)
        u32(len(exposed)); [efs(e) for e in exposed]
        u32(len(pvs)); words(pvs)
    out.append(0)
    return bytes(out), perm, logup, [1 << a[1] for a in per_air]


def wire_row_widths(proof):
    """Widths of the opened rows per query and batch opening, read off the wire bytes (needed to un-flatten)."""
    pos = [4]
    def u32():
        v = struct.unpack_from("<I", proof, pos[0])[0]; pos[0] += 4; return v
    def skip_digs():
        n = u32()
        pos[0] += 32 * n
    skip_digs(); skip_digs(); pos[0] += 32; skip_digs()
    widths = []
    for _ in range(u32()):
        q = []
        for _ in range(u32()):
            rows = []
            for _ in range(u32()):
                w = u32(); rows.append(w); pos[0] += 4 * w
            q.append(rows)
            skip_digs()
        for _ in range(u32()):
            pos[0] += 16; skip_digs()
        widths.append(q)
    return widths


def proofs(orc):
    out = []
    fib = airs.fibonacci_air()
    key = build_key(orc, [fib], 1)
    trace, pvs = airs.fibonacci_trace(32)
    out.append(("fibonacci", orc.stark_prove(key, [(0, [trace], pvs)], 1, 4, 3), False))
    snd, rcv = airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)
    key = airs.key_blob([fib, snd, rcv], log_up_pow_bits=5)
    send, recv = interaction_traces(64, 128, 2, seed=4)
    none = np.zeros(0, dtype=np.uint32)
    out.append(("logup, mixed heights", orc.stark_prove(key, [(0, [trace], pvs), (1, [send], none), (2, [recv], none)], 1, 5, 4), True))
    a0 = airs.lookup_chain_air_parts(6, 256, seed=2)
    a1 = airs.multi_receiver_air([2, 3], 5)
    key = build_key(orc, [a0, a1], 2)
    parts, pv = airs.chain_trace(a0, 256, seed=8)
    rt = airs.lookup_chain_parts_receiver_trace(a0, parts)
    out.append(("prep + cached + logup, blowup 2", orc.stark_prove(key, [(0, parts, pv), (1, [rt], none)], 2, 4, 3), True))
    return out


def test_hint_stream_product_oracle_reader_agree(orc):
    for name, proof, has_phase in proofs(orc):
        want = orc.proof_hint_stream(proof)
        got = product_stream(proof)
        assert got == want, name
        wire, perm, logup, degrees = stream_to_wire(got, wire_row_widths(proof))
        # everything up to the logup witness word is reproduced from the stream alone
        assert wire == proof[:-4], name
        witness = struct.unpack_from("<I", proof, len(proof) - 4)[0]
        assert logup == (0 if witness == 0xFFFFFFFF else witness), name
        assert (witness != 0xFFFFFFFF) == has_phase
        # air_perm_by_height: stable, descending degree (hints.rs:302-305)
        assert perm == sorted(range(len(degrees)), key=lambda i: -degrees[i]), name


def test_reference_known_answers():
    """hints.rs:455-531: Vec<Val> [1,2,3] is ONE vector; Vec<Challenge> [1,2,3] is [3] then the 12 base words. Both forms
    occur in every proof: public values, and final_poly / opened values."""
    import oracle_binding
    orc = oracle_binding.load()
    fib = airs.fibonacci_air()
    key = build_key(orc, [fib], 1)
    trace, pvs = airs.fibonacci_trace(32)
    proof = orc.stark_prove(key, [(0, [trace], pvs)], 1, 4, 3)
    st = product_stream(proof)
    assert [int(x) for x in pvs] in st                     # test_var_array's shape: the vector itself, no length
    k = st.index([int(x) for x in pvs])
    assert st[k + 1] == [0] and st[k + 2] == [0]           # then air_perm_by_height = [0] and the absent logup witness
    # final_poly: one extension element -> [1] followed by its four words in one vector (test_ext_array's shape)
    hits = [i for i in range(len(st) - 1) if st[i] == [1] and len(st[i + 1]) == 4]
    assert hits


def test_hint_stream_rejects_garbage():
    for bad in (b"\x02\x00\x00\x00", b"\x01\x00\x00\x00", b"\x01\x00\x00\x00" + b"\xff" * 40):
        rc = product_stream(bad)
        assert isinstance(rc, int) and rc != 0
