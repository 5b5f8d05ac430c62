#!/usr/bin/env python
"""One-command replay of a REFERENCE proof against this library (VERDICT r1 "pin to the reference"; PARITY.md, REPLAY.md).

    python tools/replay_reference.py --key key.b2pk --proof proof.bin [--log-blowup 1 --queries 100 --pow-bits 16]
                                     [--hash babybear|bn254] [--traces DIR] [--device 0]

Inputs, produced on a machine that can build the reference (REPLAY.md has the 30 lines of Rust that write them):
  key.b2pk    the key blob (include/b2stark.h, b2s_pk_load): little-endian u32 words written by the shim's KeyBlobWriter from the
              reference's MultiStarkProvingKey / verifying key
  proof.bin   `proof.encode_to_vec()` of the reference's Proof<SC> (crates/sdk/src/codec.rs:92-330)
  DIR         optional: air_<id>_part<k>.u32 (cached mains in order, then the common main), row-major canonical u32 words, and
              air_<id>.pv (public values); needs a GPU

Step 1 (no GPU): b2s_verify_host must ACCEPT the reference's proof. This pins, in one go, every constant and convention the
          verifier touches: field and extension, generators, Poseidon2 constants, sponge / compression / MMCS shapes, the duplex
          challenger, transcript order, domains and selectors, constraint folding, LogUp denominators, FRI folding, codec.
Step 2 (GPU, with --traces): the two proof-of-work witnesses are read out of the reference's proof (the reference returns ANY
          valid witness, this library the minimum one), handed to b2s_engine_set_pow_witnesses, the same traces are proven here,
          and the bytes are diffed. Identical bytes pin the prover side: LDE conventions, quotient, opening order, queries.
Exit status 0 = everything that was run agrees.
"""
import argparse
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def proof_sections(proof):
    """Walks the wire format far enough to name byte ranges and to find the two PoW witnesses.
    Returns (sections [(name, start, end)], fri_pow_witness, logup_pow_witness or None)."""
    pos = 0
    sections = []

    def u32():
        nonlocal pos
        if pos + 4 > len(proof):
            raise ValueError(f"proof truncated at byte {pos}")
        v = struct.unpack_from("<I", proof, pos)[0]
        pos += 4
        return v

    def skip(words):
        nonlocal pos
        pos += 4 * words
        if pos > len(proof):
            raise ValueError("proof truncated")

    def section(name, start):
        sections.append((name, start, pos))

    s = pos
    if u32() != 1:
        raise ValueError("codec version != 1")
    n_main = u32(); skip(8 * n_main); section("main commitments", s)
    s = pos
    n_after = u32(); skip(8 * n_after); section("after-challenge commitments", s)
    s = pos
    skip(8); section("quotient commitment", s)
    s = pos
    n_fri = u32(); skip(8 * n_fri); section("FRI commit-phase commitments", s)
    s = pos
    nq = u32()
    for _ in range(nq):
        for _ in range(u32()):          # rounds
            for _ in range(u32()):      # matrices
                skip(u32())
            skip(8 * u32())
        for _ in range(u32()):          # commit-phase openings
            skip(4)
            skip(8 * u32())
    section("query proofs", s)
    s = pos
    skip(u32() * 4)
    section("final polynomial", s)
    s = pos
    fri_witness = u32()
    section("FRI proof-of-work witness", s)
    section("opened values + per-AIR data", pos)
    sections[-1] = ("opened values + per-AIR data", pos, len(proof) - 5)
    sections.append(("RAP phase kind + LogUp proof-of-work witness", len(proof) - 5, len(proof)))
    logup = struct.unpack_from("<I", proof, len(proof) - 4)[0]
    return sections, fri_witness, (None if logup == 0xFFFFFFFF else logup)


def first_difference(a, b, sections):
    n = min(len(a), len(b))
    at = next((i for i in range(n) if a[i] != b[i]), None)
    if at is None:
        return None if len(a) == len(b) else (n, "length differs")
    name = next((nm for nm, s, e in sections if s <= at < e), "?")
    return at, name


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--key", required=True)
    ap.add_argument("--proof", required=True)
    ap.add_argument("--log-blowup", type=int, default=1)
    ap.add_argument("--queries", type=int, default=100)
    ap.add_argument("--pow-bits", type=int, default=16)
    ap.add_argument("--hash", default="babybear", choices=["babybear", "bn254"])
    ap.add_argument("--traces")
    ap.add_argument("--device", type=int, default=0)
    args = ap.parse_args()
    from openvm_b200 import FriParameters, VerificationError, verify_host
    key = np.fromfile(args.key, dtype="<u4")
    proof = open(args.proof, "rb").read()
    fp = FriParameters(log_blowup=args.log_blowup, num_queries=args.queries, proof_of_work_bits=args.pow_bits)
    kind = 0 if args.hash == "babybear" else 1
    try:
        sections, fri_w, logup_w = proof_sections(proof)
    except ValueError as ex:
        print(f"proof: {len(proof)} bytes; the wire format does not parse ({ex}); running the verifier anyway")
        sections, fri_w, logup_w = [], None, None
    print(f"proof: {len(proof)} bytes; FRI PoW witness {fri_w}, LogUp PoW witness {logup_w}")
    for nm, s, e in sections:
        print(f"  [{s:>9}, {e:>9})  {nm}")
    try:
        verify_host(key, proof, fp, kind)
        print("step 1: b2s_verify_host ACCEPTS the reference proof")
    except VerificationError as ex:
        print(f"step 1: b2s_verify_host REJECTS the reference proof: {ex}")
        return 1
    if not args.traces:
        print("step 2 skipped (no --traces)")
        return 0
    from openvm_b200 import AirProvingContext, B2Engine
    eng = B2Engine.new(fp, args.device)
    eng.set_hash(kind)
    dev = eng.device()
    pk = dev.transport_pk_to_device(key)
    ids = sorted({int(f.split("_")[1].split(".")[0]) for f in os.listdir(args.traces) if f.startswith("air_")})
    ctx = []
    for i in ids:
        pv = np.fromfile(os.path.join(args.traces, f"air_{i}.pv"), dtype="<u4")
        parts, k = [], 0
        while os.path.exists(os.path.join(args.traces, f"air_{i}_part{k}.u32")):
            parts.append(np.fromfile(os.path.join(args.traces, f"air_{i}_part{k}.u32"), dtype="<u4"))
            k += 1
        widths = eng.air_widths(pk, i) if hasattr(eng, "air_widths") else None
        if widths is None:
            wfile = os.path.join(args.traces, f"air_{i}.widths")
            widths = [int(x) for x in open(wfile).read().split()]
        mats = [p.reshape(-1, w) for p, w in zip(parts, widths)]
        cached, traces = [], []
        for m in mats[:-1]:
            dm = dev.transport_matrix_to_device(m)
            cached.append(dev.commit([dm])[1])
            traces.append(dm)
        ctx.append(AirProvingContext(i, dev.transport_matrix_to_device(mats[-1]), pv, tuple(cached), tuple(traces)))
    eng.set_pow_witnesses(logup_w, fri_w)
    mine = eng.prove(pk, ctx)
    diff = first_difference(mine, proof, sections)
    if diff is None:
        print("step 2: the proof made here from the same traces and witnesses is BYTE-IDENTICAL to the reference's")
        return 0
    print(f"step 2: proofs differ first at byte {diff[0]} ({diff[1]}); this library's proof: {len(mine)} bytes")
    return 2


if __name__ == "__main__":
    sys.exit(main())
