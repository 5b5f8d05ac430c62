"""Self-checks of the 'parity unpinned' constants: they are recalled from the un-vendored
dependencies (Plonky3 @539bbc8, zkhash @bb476b9), so each gets the strongest check available in-tree."""
import numpy as np

import pyref

P = pyref.P


def test_generator_is_primitive_root():
    # p - 1 = 2^27 * 3 * 5
    assert (P - 1) == (1 << 27) * 15
    for q in (2, 3, 5):
        assert pow(31, (P - 1) // q, P) != 1


def test_two_adic_generator_order(orc):
    g27 = orc.two_adic_generator(27)
    assert g27 == 0x1A427A41
    assert pow(g27, 1 << 26, P) == P - 1
    for bits in range(0, 28):
        g = orc.two_adic_generator(bits)
        assert g == pyref.two_adic_generator(bits)
        assert pow(g, 1 << bits, P) == 1
        if bits:
            assert pow(g, 1 << (bits - 1), P) == P - 1


def test_rc16_first_row_matches_published_words(orc):
    # leading words of zkhash poseidon2_instance_babybear.rs RC16 (recalled); Grain LFSR regenerates the table
    want = [0x69CBB6AF, 0x46AD93F9, 0x60A00F4E, 0x6B1297CD, 0x23189AFE, 0x732E7BEF, 0x72C246DE, 0x2C941900,
            0x0557EEDE, 0x1580496F, 0x3A3EA77B, 0x54F3F271, 0x0F49B029, 0x47872FE1, 0x221E2E36, 0x1AB7202E]
    c = orc.poseidon2_constants()
    assert [int(x) for x in c["ext_init"][0]] == want
    assert np.array_equal(c["ext_init"], np.array(pyref.EXT_INIT, dtype=np.uint32))
    assert np.array_equal(c["partial"], np.array(pyref.PARTIAL, dtype=np.uint32))
    assert np.array_equal(c["ext_term"], np.array(pyref.EXT_TERM, dtype=np.uint32))
    assert all(0 < int(x) < P for x in c["partial"])


def test_internal_diag(orc):
    d = [int(x) for x in orc.poseidon2_constants()["diag"]]
    # first three lanes are pinned in-tree (linearlayer.cuh:84-94): -2, 1, 2
    assert d[:3] == [P - 2, 1, 2]
    assert d == pyref.DIAG
    # inverse powers of two really are inverses
    assert d[3] * 2 % P == 1 and d[9] * 256 % P == 1 and d[12] * (1 << 27) % P == 1
    # 1 + diag(d) must be invertible for the layer to be a permutation: det = prod(d) * (1 + sum(1/d))
    s = sum(pyref.inv(x) for x in d) % P
    assert (1 + s) % P != 0 and all(x != 0 for x in d)


def test_generated_header_is_current():
    import os, subprocess, sys
    root = os.path.join(os.path.dirname(__file__), "..")
    path = os.path.join(root, "openvm_b200", "csrc", "poseidon2_constants.cuh")
    before = open(path).read()
    subprocess.check_call([sys.executable, os.path.join(root, "tools", "gen_constants.py")], stdout=subprocess.DEVNULL)
    assert open(path).read() == before
    R = (1 << 32) % P
    assert ("0x%08xu" % (0x69CBB6AF * R % P)) in before
