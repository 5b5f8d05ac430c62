"""Committed fixtures (tests/golden/, made by tests/golden/make_golden.py): proof bytes and commitment roots of the
reference's own small fixture inputs under both configurations, plus the published known answers. They are frozen oracle
outputs, not reference outputs (the reference's prover cannot run here) -- their job is to stop the oracle
and the product from drifting together. CPU: the oracle still reproduces them. GPU: so does the product, through the C ABI."""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLDEN)
INDEX = json.load(open(os.path.join(GOLDEN, "index.json")))


def read(fn):
    return open(os.path.join(GOLDEN, fn), "rb").read()


def test_index_matches_files():
    assert len(INDEX["proofs"]) == 4
    for fn, meta in INDEX["proofs"].items():
        data = read(fn)
        assert len(data) == meta["bytes"] and hashlib.sha256(data).hexdigest() == meta["sha256"]


def test_oracle_reproduces_the_fixtures(orc):
    import make_golden
    files, index = make_golden.generate(orc)
    assert index == INDEX
    for fn, data in files.items():
        assert data == read(fn), fn


def test_published_known_answers():
    ka = INDEX["known_answers"]
    assert ka["poseidon2_bn254_of_0_1_2 (published)"][0] == "0xbb61d24daca55eebcb1929a82650f328134334da98ea4f847f760054f4a3033"
    assert ka["rc16_first_row (published leading words)"][:4] == ["0x69cbb6af", "0x46ad93f9", "0x60a00f4e", "0x6b1297cd"]


def test_fixture_proofs_verify_and_cross_reject(orc):
    import make_golden
    for kind in (0, 1):
        orc.set_hash(kind)
        try:
            for name, (key, _, b, q, pw) in make_golden.cases(orc).items():
                for tag, k2 in (("inner", 0), ("root", 1)):
                    verdict = orc.stark_verify(key, read(f"{name}_{tag}.proof"), b, q, pw)
                    assert (verdict is None) == (k2 == kind), (name, tag, kind, verdict)
        finally:
            orc.set_hash(0)


@pytest.mark.gpu
def test_product_reproduces_the_fixtures(orc):
    import make_golden
    from openvm_b200 import B2Engine, FriParameters, VerificationError
    from test_gpu_prove import gpu_prove
    from openvm_b200 import airs
    objs = {"fib32": [airs.fibonacci_air()],
            "fib8_logup": [airs.fibonacci_air(), airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)]}
    for kind, tag in ((0, "inner"), (1, "root")):
        for name, (key, inputs, b, q, pw) in make_golden.cases(orc).items():
            eng = B2Engine.new(FriParameters(log_blowup=b, num_queries=q, proof_of_work_bits=pw))
            try:
                eng.set_hash(kind)
                assert gpu_prove(eng, key, inputs, objs[name]) == read(f"{name}_{tag}.proof"), (name, tag)
                pk = eng.device().transport_pk_to_device(key)
                eng.verify(pk, read(f"{name}_{tag}.proof"))
                other = read(f"{name}_{'root' if kind == 0 else 'inner'}.proof")
                with pytest.raises(VerificationError):
                    eng.verify(pk, other)
            finally:
                eng.close()
        eng = B2Engine.new(FriParameters(log_blowup=1))
        try:
            eng.set_hash(kind)
            dev = eng.device()
            d = dev.alloc_matrix(1 << 8, 40)
            dev.fill_synthetic(d, 0xB200000000000001)
            root, _ = dev.commit([d])
            assert [int(x) for x in root] == INDEX["commit"][f"synthetic_2^8x40_b1_{tag}"]
        finally:
            eng.close()


def test_product_verifier_accepts_the_fixtures_on_the_host(orc):
    """b2s_verify_host (no device) on the frozen proofs: accepted under their configuration, rejected under the other."""
    import make_golden
    from openvm_b200 import FriParameters, VerificationError, verify_host
    for name, (key, _, b, q, pw) in make_golden.cases(orc).items():
        fp = FriParameters(log_blowup=b, num_queries=q, proof_of_work_bits=pw)
        for tag, kind in (("inner", 0), ("root", 1)):
            verify_host(key, read(f"{name}_{tag}.proof"), fp, kind)
            with pytest.raises(VerificationError):
                verify_host(key, read(f"{name}_{tag}.proof"), fp, 1 - kind)
