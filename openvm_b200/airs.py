"""Key-blob writer (format: include/b2stark.h, b2s_pk_load) and sample AIRs with trace generators.

Host-side stand-in for the part of the reference's keygen that this repo cannot run (symbolic AIR building lives in
Rust: `circuit.keygen(&engine)`, crates/vm/src/arch/vm.rs:381); used by the tests and by bench.py's proof leg.

A tiny symbolic builder produces the SymbolicExpressionDag node list the way the reference's keygen does
(topological order, constraints listed by node index). The AIRs mirror the reference's own fixtures:
  * FibonacciAir(0, 1, n): 2 columns, 3 public values -- openvm_stark_sdk::dummy_airs::fib_air as used in
    extensions/native/recursion/src/tests.rs:40,109,118 and benchmarks/prove/src/bin/verify_fibair.rs:38-39
  * ChainAir: SURVEY.md 8(d) "chain AIR" (degree-3 constraints -> quotient degree 2), with optional preprocessed
    and cached-main partitions so every kind of opening round is exercised.
"""
import numpy as np

P = 2013265921
(N_CONST, N_PREP, N_MAIN, N_PERM, N_PUBLIC, N_CHALLENGE, N_EXPOSED, N_FIRST, N_LAST, N_TRANS, N_ADD, N_SUB, N_NEG,
 N_MUL) = range(14)


class Dag:
    def __init__(self):
        self.nodes = []
        self.index = {}
        self.constraints = []

    def _node(self, kind, a=0, b=0, c=0):
        key = (kind, a, b, c)
        if key not in self.index:
            self.index[key] = len(self.nodes)
            self.nodes.append(key)
        return Expr(self, self.index[key])

    def const(self, v): return self._node(N_CONST, v % P)
    def prep(self, col, offset=0): return self._node(N_PREP, col, offset)
    def main(self, col, offset=0, part=0): return self._node(N_MAIN, col, offset, part)
    def public(self, i): return self._node(N_PUBLIC, i)
    def perm(self, col, offset=0): return self._node(N_PERM, col, offset)
    def challenge(self, i): return self._node(N_CHALLENGE, i)
    def exposed(self, i): return self._node(N_EXPOSED, i)
    def is_first(self): return self._node(N_FIRST)
    def is_last(self): return self._node(N_LAST)
    def is_transition(self): return self._node(N_TRANS)

    def assert_zero(self, e):
        self.constraints.append(e.i)


class Expr:
    def __init__(self, dag, i):
        self.dag, self.i = dag, i

    def _lift(self, o):
        return o if isinstance(o, Expr) else self.dag.const(o)

    def __add__(self, o): return self.dag._node(N_ADD, self.i, self._lift(o).i)
    def __sub__(self, o): return self.dag._node(N_SUB, self.i, self._lift(o).i)
    def __mul__(self, o): return self.dag._node(N_MUL, self.i, self._lift(o).i)
    def __neg__(self): return self.dag._node(N_NEG, self.i)


class Air:
    """Key-side description of one AIR."""

    def __init__(self, dag, width_common, num_pv, log_qd, cached_widths=(), prep=None):
        self.dag, self.width_common, self.num_pv, self.log_qd = dag, width_common, num_pv, log_qd
        self.cached_widths = list(cached_widths)
        self.prep = prep            # (height, width) uint32 array or None
        self.prep_commit = None     # 8 words, filled by the key builder
        self.interactions = []      # (bus, type 0 send / 1 receive, count Expr, [field Exprs])
        self.chunks = []            # lists of interaction indices, one reciprocal-sum column each

    def add_interaction(self, bus, fields, count, is_send=True):
        self.interactions.append((bus, 0 if is_send else 1, count, list(fields)))

    def finalize_logup(self, max_chunk=1):
        """Append the FRI-LogUp constraints to the DAG (what the reference's keygen does through
        `eval_fri_log_up_phase`; formula recalled, PARITY UNPINNED -- DESIGN.md):
          denom_i = alpha + sum_j beta^j * field_ij + beta^len_i * (bus_i + 1)
          per chunk c:  perm_c * prod_i denom_i = sum_i (+-count_i) * prod_{k != i} denom_k
          phi (last column) is the running sum of sum_c perm_c; its final value is the exposed cumulative sum."""
        if not self.interactions:
            return self
        d = self.dag
        alpha, beta = d.challenge(0), d.challenge(1)
        max_len = max(len(f) for _, _, _, f in self.interactions)
        betas = [d.const(1)]
        for _ in range(max_len):
            betas.append(betas[-1] * beta)
        denoms, counts = [], []
        for bus, typ, count, fields in self.interactions:
            den = alpha
            for j, f in enumerate(fields):
                den = den + (betas[j] * f if j else f)
            den = den + betas[len(fields)] * (bus + 1)
            denoms.append(den)
            counts.append(-count if typ else count)
        idx = list(range(len(self.interactions)))
        self.chunks = [idx[i:i + max_chunk] for i in range(0, len(idx), max_chunk)]
        nch = len(self.chunks)
        for c, chunk in enumerate(self.chunks):
            prod = denoms[chunk[0]]
            for i in chunk[1:]:
                prod = prod * denoms[i]
            rhs = None
            for i in chunk:
                term = counts[i]
                for k in chunk:
                    if k != i:
                        term = term * denoms[k]
                rhs = term if rhs is None else rhs + term
            d.assert_zero(d.perm(c) * prod - rhs)
        phi, nphi = d.perm(nch), d.perm(nch, 1)
        s0, s1 = d.perm(0), d.perm(0, 1)
        for c in range(1, nch):
            s0, s1 = s0 + d.perm(c), s1 + d.perm(c, 1)
        d.assert_zero(d.is_first() * (phi - s0))
        d.assert_zero(d.is_transition() * (nphi - phi - s1))
        d.assert_zero(d.is_last() * (phi - d.exposed(0)))
        return self


def key_blob(airs, pre_hash=None, log_up_pow_bits=0, trace_height_constraints=None):
    """airs: list of Air (prep_commit set where prep is not None). Returns the uint32 word array.
    trace_height_constraints: list of (coefficients[one per AIR], threshold) -- the verifying key's LinearConstraint list
    (extensions/native/recursion/src/types.rs:70); given, the blob is version 3 and carries them after the AIRs."""
    w = [0x4B503242, 2 if trace_height_constraints is None else 3]
    w += list(pre_hash if pre_hash is not None else range(101, 109))
    w += [log_up_pow_bits, len(airs)]
    for a in airs:
        w += [0 if a.prep is None else a.prep.shape[1], len(a.cached_widths)] + a.cached_widths
        nch = len(a.chunks)
        w += [a.width_common, nch + 1 if nch else 0, a.log_qd, a.num_pv, 1 if nch else 0]
        if a.prep is not None:
            assert a.prep_commit is not None
            w += [int(x) for x in a.prep_commit]
            w += [a.prep.shape[0].bit_length() - 1]
            w += [int(x) for x in a.prep.reshape(-1)]
        w += [len(a.dag.nodes)]
        for n in a.dag.nodes:
            w += list(n)
        w += [len(a.dag.constraints)] + a.dag.constraints
        w += [len(a.interactions)]
        for bus, typ, count, fields in a.interactions:
            w += [bus, typ, count.i, len(fields)] + [f.i for f in fields]
        w += [nch]
        for ch in a.chunks:
            w += [len(ch)] + ch
    if trace_height_constraints is not None:
        w += [len(trace_height_constraints)]
        for coeffs, threshold in trace_height_constraints:
            assert len(coeffs) == len(airs)
            w += [int(c) for c in coeffs] + [int(threshold)]
    return np.array(w, dtype=np.uint32)


# ---------------------------------------------------------------------------------------------------------------
def fibonacci_air():
    d = Dag()
    left, right = d.main(0), d.main(1)
    nleft, nright = d.main(0, 1), d.main(1, 1)
    a, b, x = d.public(0), d.public(1), d.public(2)
    d.assert_zero(d.is_first() * (left - a))
    d.assert_zero(d.is_first() * (right - b))
    d.assert_zero(d.is_transition() * (nleft - right))
    d.assert_zero(d.is_transition() * (nright - (left + right)))
    d.assert_zero(d.is_last() * (right - x))
    return Air(d, width_common=2, num_pv=3, log_qd=0)


def fibonacci_trace(n, a=0, b=1):
    t = np.zeros((n, 2), dtype=np.uint32)
    for i in range(n):
        t[i] = (a, b)
        a, b = b, (a + b) % P
    return t, np.array([t[0, 0], t[0, 1], t[n - 1, 1]], dtype=np.uint32)


def chain_air(width, with_prep=False, cached_width=0, height=None, seed=0, degree=3):
    """Columns x_0..x_{W-1} of the common main: x_{j+2} = x_j * x_{j+1} + x_j (degree 2), a degree-3 booleanity
    style constraint x_0 * (x_1 * x_2 - x_3 + x_0) = 0 ... kept satisfiable by construction: instead we use
    y := x_j * x_{j+1} * x_{j+2} - x_{j+3}' terms only on columns defined that way. Layout:
      common main: x_0, x_1 free per row; x_{j+2} = x_j * x_{j+1} + x_j for j + 2 < W - 1;
                   last column c = x_0 * x_1 * x_2 (degree 3)
      transition: next.x_0 = local.x_0 + 1 ; first row: x_0 = pv_0 ; last row: x_1 = pv_1
      preprocessed (optional, width 2): p_0 = row index, p_1 = seed-derived; constraint x_1 = p_0 * p_1 + pv_0
      cached main (optional): k columns, each = x_0 + column index (degree 1)
      degree 5 (quotient degree 4, needs log_blowup >= 2): the same traces, the last-column constraint multiplied by
      x_3 * x_4 and a second one (x_2 - x_0 * x_1 - x_0) * x_1 * x_3 * x_4
    """
    assert width >= 5 and degree in (3, 5)
    d = Dag()
    x = [d.main(j, 0, 1 if cached_width else 0) for j in range(width)]
    nx0 = d.main(0, 1, 1 if cached_width else 0)
    for j in range(width - 3):
        d.assert_zero(x[j + 2] - (x[j] * x[j + 1] + x[j]))
    if degree == 3:
        d.assert_zero(x[width - 1] - x[0] * x[1] * x[2])
    else:
        d.assert_zero((x[width - 1] - x[0] * x[1] * x[2]) * x[3] * x[4])
        d.assert_zero((x[2] - x[0] * x[1] - x[0]) * x[1] * x[3] * x[4])
    d.assert_zero(d.is_transition() * (nx0 - x[0] - 1))
    d.assert_zero(d.is_first() * (x[0] - d.public(0)))
    d.assert_zero(d.is_last() * (x[1] - d.public(1)))
    prep = None
    if with_prep:
        assert height is not None
        rng = np.random.default_rng(1000 + seed)
        prep = np.stack([np.arange(height, dtype=np.uint64) % P,
                         rng.integers(0, P, size=height, dtype=np.uint64)], axis=1).astype(np.uint32)
        d.assert_zero(x[1] - (d.prep(0) * d.prep(1) + d.public(0)))
        # a rotated preprocessed access
        d.assert_zero(d.is_transition() * (d.prep(0, 1) - d.prep(0) - 1))
    for k in range(cached_width):
        d.assert_zero(d.main(k, 0, 0) - x[0] - k)
    return Air(d, width_common=width, num_pv=2, log_qd=1 if degree == 3 else 2,
               cached_widths=[cached_width] if cached_width else [], prep=prep)


_LCG_A, _LCG_C = 6364136223846793005, 1442695040888963407


def _hash_pv0(seed):
    return ((seed * _LCG_A + _LCG_C) % (1 << 64) >> 24) % P


def chain_trace(air, height, seed=0, hashed=False):
    """Returns ([cached parts..., common], public values) satisfying chain_air. `hashed`: the free values come from the
    arithmetic formula chain_trace_device uses on the GPU instead of numpy's generator (same trace on both sides)."""
    W = air.width_common
    rng = np.random.default_rng(seed)
    pv0 = _hash_pv0(seed) if hashed else int(rng.integers(0, P))
    p = np.uint64(P)
    t = np.zeros((height, W), dtype=np.uint64)
    t[:, 0] = (np.uint64(pv0) + np.arange(height, dtype=np.uint64)) % p
    if air.prep is not None:
        t[:, 1] = (air.prep[:, 0].astype(np.uint64) * air.prep[:, 1].astype(np.uint64) + np.uint64(pv0)) % p
    elif hashed:
        with np.errstate(over="ignore"):
            x = (np.arange(height, dtype=np.uint64) + np.uint64(seed * 1000003 % (1 << 62))) * np.uint64(_LCG_A) + np.uint64(_LCG_C)
        t[:, 1] = ((x >> np.uint64(20)) & np.uint64((1 << 40) - 1)) % p
    else:
        t[:, 1] = rng.integers(0, P, size=height, dtype=np.uint64)
    for j in range(W - 3):
        t[:, j + 2] = (t[:, j] * t[:, j + 1] + t[:, j]) % p
    t[:, W - 1] = ((t[:, 0] * t[:, 1]) % p * t[:, 2]) % p
    common = t.astype(np.uint32)
    parts = []
    for cw in air.cached_widths:
        parts.append(np.stack([(t[:, 0] + np.uint64(k)) % p for k in range(cw)], axis=1).astype(np.uint32))
    parts.append(common)
    return parts, np.array([pv0, int(common[height - 1, 1])], dtype=np.uint32)


def chain_trace_device(air, height, seed, device):
    """chain_trace(air, height, seed, hashed=True) computed ON THE GPU with torch ops on torch's current stream: returns
    (columns, public values) where `columns` is a (width, height) int32 CUDA tensor = the column-major, Montgomery-form
    matrix a device tracegen would leave in HBM (the reference's DeviceMatrix, crates/vm/src/system/cuda/program.rs:60,
    119-125), ready for b2s_matrix_adopt_device without any copy. Only for AIRs with a single common-main partition."""
    import torch
    assert air.prep is None and not air.cached_widths
    W = air.width_common
    pv0 = _hash_pv0(seed)
    r = torch.arange(height, dtype=torch.int64, device=device)
    t = torch.empty((W, height), dtype=torch.int64, device=device)
    t[0] = (r + pv0) % P
    x = (r + (seed * 1000003 % (1 << 62))) * _signed64(_LCG_A) + _signed64(_LCG_C)     # wraps mod 2^64 like the host formula
    t[1] = ((x >> 20) & ((1 << 40) - 1)) % P
    for j in range(W - 3):
        t[j + 2] = (t[j] * t[j + 1] + t[j]) % P
    t[W - 1] = ((t[0] * t[1]) % P * t[2]) % P
    # the second public value is the last row's x_1: recomputed here on the host, so that nothing waits on the device
    xl = ((height - 1 + seed * 1000003 % (1 << 62)) * _LCG_A + _LCG_C) % (1 << 64)
    pv1 = ((xl >> 20) & ((1 << 40) - 1)) % P
    monty = ((t << 32) % P).to(torch.int32).contiguous()
    return monty, np.array([pv0, pv1], dtype=np.uint32)


def _signed64(x):
    x %= 1 << 64
    return x - (1 << 64) if x >= 1 << 63 else x


def dummy_interaction_air(n_fields, is_send, bus=0, log_qd=1):
    """DummyInteractionAir of the reference's tests (openvm_stark_sdk::dummy_airs::interaction, used at
    extensions/native/recursion/src/tests.rs:57-68,170-181): columns [count, field_0..]; one interaction, no other
    constraints."""
    d = Dag()
    count = d.main(0)
    fields = [d.main(1 + j) for j in range(n_fields)]
    a = Air(d, width_common=1 + n_fields, num_pv=0, log_qd=log_qd)
    a.add_interaction(bus, fields, count, is_send)
    return a.finalize_logup(1)


def lookup_chain_air(width, bus=3):
    """chain_air whose rows are also sent on a bus: (x_0, x_1) with multiplicity 1 and (x_2) on bus+1 with multiplicity
    x_3's low bit is not needed -- multiplicity 1; two interactions in one chunk (degree 3)."""
    a = chain_air(width)
    d = a.dag
    a.add_interaction(bus, [d.main(0), d.main(1)], d.const(1), True)
    a.add_interaction(bus + 1, [d.main(2)], d.const(1), True)
    return a.finalize_logup(2)


def lookup_chain_receiver(bus=3):
    """receives what lookup_chain_air sends: columns [m, a, b, c]: receive (a, b) on `bus` and (c) on bus+1, count m."""
    d = Dag()
    m, a_, b_, c_ = (d.main(j) for j in range(4))
    a = Air(d, width_common=4, num_pv=0, log_qd=1)
    a.add_interaction(bus, [a_, b_], m, False)
    a.add_interaction(bus + 1, [c_], m, False)
    return a.finalize_logup(2)


def build_key(air_list, commit_fn, **kw):
    """Fill in the preprocessed commitments with `commit_fn(matrix) -> 8 canonical words` and return the key blob."""
    for a in air_list:
        if a.prep is not None:
            a.prep_commit = commit_fn(a.prep)
    return key_blob(air_list, **kw)


# (rows, main columns, after-challenge EF columns = perm_cols / 4) of every AIR present in segment 0 of the reference's example
# run, group ecrecover_program of crates/prof/data/metrics.example.json (the table SURVEY.md 8d names), in AIR order
_ECRECOVER_SEGMENT = [
    (16384, 10, 2), (2, 4, 3), (4096, 20, 3), (4096, 32, 5), (256, 11, 6), (128, 13, 6), (32768, 17, 6), (16384, 25, 6),
    (8192, 41, 6), (2048, 543, 207), (1024, 619, 212), (128, 3164, 322), (4096, 166, 14), (16, 261, 79), (8, 199, 48),
    (256, 26, 9), (8, 57, 26), (8, 39, 25), (4096, 31, 20), (524288, 1, 2), (4096, 21, 7), (8192, 28, 9), (4096, 18, 11),
    (32768, 32, 14), (16384, 26, 12), (4096, 35, 19), (131072, 40, 18), (16384, 53, 13), (2048, 37, 10), (131072, 36, 20),
    (65536, 2, 2), (64, 6, 3), (4096, 300, 2), (262144, 1, 2)]


def segment_shapes(scale_log=0, legacy=False):
    """Shape replay of an RV32IM app segment: (log_height, main width, after-challenge EF columns) per AIR = the per-AIR table
    of the reference's example metrics (34 AIRs with rows in segment 0: BaseAlu 131072 x 36 (+80 perm words), LoadStore
    131072 x 40 (+72), RangeTuple 524288 x 1 (+8), KeccakVm 128 x 3164 (+1288) ...). scale_log = 3 makes the ALU / load-store
    chips 2^20 rows. Main widths below the chain AIR's minimum of 5 are raised to 5.
    legacy: the ten-AIR list of round 1 (kept for comparisons with profiles/r1*)."""
    if legacy:
        base = [(19, 5), (17, 36), (17, 40), (16, 28), (16, 20), (15, 32), (14, 66), (13, 200), (12, 16), (10, 8)]
        return [(max(4, h + scale_log), w, None) for h, w in base]
    return [(max(1, rows.bit_length() - 1 + scale_log), max(5, w), perm) for rows, w, perm in _ECRECOVER_SEGMENT]


def leaf_shapes(scale_log=0):
    """Shape replay of a leaf-aggregation proof (15 AIRs, heights up to 2^22, ~446 M main cells,
    quotient degree up to 4, log_blowup 2): (log_height, main width, constraint degree) per AIR."""
    base = [(22, 34, 3), (21, 48, 5), (21, 24, 3), (20, 48, 3), (20, 30, 5), (19, 40, 3), (19, 28, 3), (18, 36, 5),
            (18, 20, 3), (17, 66, 3), (16, 200, 3), (15, 32, 3), (14, 16, 3), (12, 8, 3), (10, 8, 3)]
    return [(max(4, h + scale_log), w, dg) for h, w, dg in base]


def lookup_chain_air_parts(width, height, seed=0, bus=5, cached_width=2):
    """chain_air with a preprocessed trace and a cached partition whose interactions read every kind of column:
    send (x_0, x_1) once per row on `bus`; send (cached_0, prep_0, next x_0) with multiplicity prep_0 (the row index)
    on bus + 1. One interaction per chunk -> two reciprocal-sum columns."""
    a = chain_air(width, with_prep=True, cached_width=cached_width, height=height, seed=seed)
    d = a.dag
    a.add_interaction(bus, [d.main(0, 0, 1), d.main(1, 0, 1)], d.const(1), True)
    a.add_interaction(bus + 1, [d.main(0, 0, 0), d.prep(0), d.main(0, 1, 1)], d.prep(0), True)
    return a.finalize_logup(1)


def multi_receiver_air(field_lens, bus0, max_chunk=1, is_send=False, log_qd=1):
    """columns [m_0, fields_0.., m_1, fields_1.., ...]: interaction i receives (or sends) its fields with multiplicity
    m_i on bus0 + i. `max_chunk` interactions share one reciprocal-sum column: constraint degree 1 + max_chunk, so
    three per chunk need quotient degree 4 (log_qd 2)."""
    d = Dag()
    col = 0
    a = Air(d, width_common=sum(field_lens) + len(field_lens), num_pv=0, log_qd=log_qd)
    for i, ln in enumerate(field_lens):
        m = d.main(col)
        a.add_interaction(bus0 + i, [d.main(col + 1 + j) for j in range(ln)], m, is_send)
        col += 1 + ln
    return a.finalize_logup(max_chunk)


def lookup_chain_parts_receiver_trace(air, parts):
    """receiver rows for lookup_chain_air_parts: the same messages, in reverse row order."""
    cached, common = parts[0], parts[-1]
    n = common.shape[0]
    prep0 = air.prep[:, 0]
    t = np.zeros((n, 7), dtype=np.uint32)
    t[:, 0] = 1
    t[:, 1:3] = common[:, 0:2]
    t[:, 3] = prep0
    t[:, 4] = cached[:, 0]
    t[:, 5] = prep0
    t[:, 6] = np.roll(common[:, 0], -1)
    return t[::-1].copy()


def segment_air(width, degree=3, perm_cols=None):
    """One chip of the shape-replay segment: chain_air plus LogUp traffic shaped like an OpenVM chip's (BaseAlu:
    36 main columns and 20 extension-field permutation columns in crates/prof/data/metrics.example.json). Chunk c holds
    two interactions on bus c: send (x_j, x_{j+1}) of this row and receive the same pair of the NEXT row, so the bus
    balances over the cyclic trace while every row's reciprocal sum is non-trivial. Degree 3 -> quotient degree 2.
    perm_cols: after-challenge EF columns wanted (chunks + the running sum); default: the round-1 heuristic."""
    a = chain_air(width, degree=degree)
    d = a.dag
    if perm_cols is not None:
        k = max(1, perm_cols - 1)
    else:
        k = max(1, width // 2) if width <= 100 else width // 10
    for c in range(k):
        j = c % (width - 1)
        a.add_interaction(c, [d.main(j), d.main(j + 1)], d.const(1), True)
        a.add_interaction(c, [d.main(j, 1), d.main(j + 1, 1)], d.const(1), False)
    return a.finalize_logup(2)
