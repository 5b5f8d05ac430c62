"""tools/replay_reference.py (REPLAY.md) on an oracle-made proof: the command a maintainer with a reference proof runs."""
import os
import subprocess
import sys

import numpy as np
import pytest

from openvm_b200 import airs
from test_oracle_stark import interaction_traces

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NONE = np.zeros(0, dtype=np.uint32)


def _fixture(orc, tmp_path):
    air_objs = [airs.fibonacci_air(), airs.dummy_interaction_air(2, True), airs.dummy_interaction_air(2, False)]
    key = airs.key_blob(air_objs, log_up_pow_bits=3)
    send, recv = interaction_traces(16, 32, 2, seed=4)
    t0, pv0 = airs.fibonacci_trace(32)
    inputs = [(0, [t0], pv0), (1, [send], NONE), (2, [recv], NONE)]
    proof = orc.stark_prove(key, inputs, 1, 5, 4)
    key.astype("<u4").tofile(tmp_path / "key.b2pk")
    (tmp_path / "proof.bin").write_bytes(proof)
    tdir = tmp_path / "traces"
    tdir.mkdir()
    for i, parts, pv in inputs:
        for k, m in enumerate(parts):
            np.ascontiguousarray(m, dtype="<u4").tofile(tdir / f"air_{i}_part{k}.u32")
        (tdir / f"air_{i}.widths").write_text(" ".join(str(m.shape[1]) for m in parts))
        np.asarray(pv, dtype="<u4").tofile(tdir / f"air_{i}.pv")
    return proof


def _run(tmp_path, *extra):
    cmd = [sys.executable, os.path.join(ROOT, "tools", "replay_reference.py"), "--key", str(tmp_path / "key.b2pk"), "--proof",
           str(tmp_path / "proof.bin"), "--log-blowup", "1", "--queries", "5", "--pow-bits", "4", *extra]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600)


def test_replay_step1_on_cpu(orc, tmp_path):
    proof = _fixture(orc, tmp_path)
    r = _run(tmp_path)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ACCEPTS" in r.stdout and "step 2 skipped" in r.stdout
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import replay_reference as rr
    sections, fri_w, logup_w = rr.proof_sections(proof)
    assert sections[0][1] == 0 and sections[-1][2] == len(proof)
    assert all(a[2] == b[1] for a, b in zip(sections, sections[1:]))      # the sections tile the proof
    assert logup_w is not None
    # a tampered proof is rejected and the tool says so
    for at in (12, 40):                      # a digest word; a length field (the section walk fails, the verifier still runs)
        bad = bytearray(proof); bad[at] ^= 1
        (tmp_path / "proof.bin").write_bytes(bytes(bad))
        r = _run(tmp_path)
        assert r.returncode == 1 and "REJECTS" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_replay_step2_reproves_byte_identically(orc, tmp_path):
    _fixture(orc, tmp_path)
    r = _run(tmp_path, "--traces", str(tmp_path / "traces"))
    assert r.returncode == 0, r.stdout + r.stderr
    assert "BYTE-IDENTICAL" in r.stdout
