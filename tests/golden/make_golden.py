"""Regenerates the fixtures in this directory from the CPU oracle:  python tests/golden/make_golden.py

The reference's prover lives in un-vendored Rust crates and cannot run here, so these are NOT outputs of the
reference: they freeze the oracle's outputs on the reference's own fixture inputs (FibonacciChip::new(0, 1, n),
extensions/native/recursion/src/tests.rs:40,109; DummyInteraction tables :57-68), so that the oracle and the product cannot
drift together unnoticed. The published known answers that DO pin something (Poseidon2 over BN254; the leading round
constants) sit in known_answers.json as well. Inputs are small enough that the files stay a few kilobytes each."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from openvm_b200 import airs  # noqa: E402
import oracle_binding  # noqa: E402
import pyref  # noqa: E402


def cases(orc):
    """name -> (hash kind, key words, prover inputs, log_blowup, queries, pow bits)"""
    from test_oracle_stark import interaction_traces
    fib = airs.fibonacci_air()
    t32, pv32 = airs.fibonacci_trace(32)
    out = {"fib32": (airs.key_blob([fib]), [(0, [t32], pv32)], 1, 4, 3)}
    snd, rcv = airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)
    send, recv = interaction_traces(16, 32, 2, seed=4)
    t8, pv8 = airs.fibonacci_trace(8)
    none = np.zeros(0, dtype=np.uint32)
    out["fib8_logup"] = (airs.key_blob([fib, snd, rcv], log_up_pow_bits=3), [(0, [t8], pv8), (1, [send], none), (2, [recv], none)], 1, 3, 2)
    return out


def generate(orc):
    files, index = {}, {"proofs": {}, "commit": {}, "known_answers": {}}
    for kind, tag in ((0, "inner"), (1, "root")):
        orc.set_hash(kind)
        try:
            for name, (key, inputs, b, q, pw) in cases(orc).items():
                proof = orc.stark_prove(key, inputs, b, q, pw)
                assert orc.stark_verify(key, proof, b, q, pw) is None
                fn = f"{name}_{tag}.proof"
                files[fn] = proof
                index["proofs"][fn] = {"log_blowup": b, "num_queries": q, "pow_bits": pw, "hash_kind": kind,
                                       "sha256": hashlib.sha256(proof).hexdigest(), "bytes": len(proof)}
            m = orc.fill_synthetic(1 << 8, 40, 0xB200000000000001)
            index["commit"][f"synthetic_2^8x40_b1_{tag}"] = [int(x) for x in orc.commit([m], 1)]
        finally:
            orc.set_hash(0)
    index["known_answers"]["poseidon2_babybear16_of_0..15"] = [int(x) for x in orc.permute(np.arange(16, dtype=np.uint32))[0]]
    index["known_answers"]["poseidon2_bn254_of_0_1_2 (published)"] = [hex(x) for x in pyref.bn_permute([0, 1, 2])]
    index["known_answers"]["rc16_first_row (published leading words)"] = [hex(x) for x in pyref.EXT_INIT[0]]
    return files, index


if __name__ == "__main__":
    orc = oracle_binding.load()
    files, index = generate(orc)
    for fn, data in files.items():
        with open(os.path.join(HERE, fn), "wb") as f:
            f.write(data)
    with open(os.path.join(HERE, "index.json"), "w") as f:
        json.dump(index, f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote", len(files), "proofs and index.json")
