"""The C-ABI library loads on a CPU-only box and exports every symbol include/b2stark.h declares.
No compute call is made here (there is no GPU); constructing an engine must fail loudly, not fall back."""
import ctypes
import os
import re

import pytest

ROOT = os.path.join(os.path.dirname(__file__), "..")


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b2stark.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2s_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    from openvm_b200 import build, _lib
    build.build()
    return _lib.load()


def test_every_declared_symbol_is_exported_and_bound(lib):
    from openvm_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in b2stark.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(names)


def test_library_is_sm100a_only(lib):
    import subprocess
    from openvm_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert not re.search(r"sm_(?!100a)\d+", out)


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from openvm_b200 import B2Engine, FriParameters, B2StarkError
    with pytest.raises(B2StarkError) as ei:
        B2Engine.new(FriParameters(log_blowup=1))
    assert ei.value.status == -4  # B2S_ERR_NO_DEVICE


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "openvm_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle_binding" not in src and "liboracle" not in src and "oracle/" not in src, f


def test_quotient_ptx_assembles_for_sm_100a(tmp_path):
    """The quotient kernels are generated as PTX from the constraint DAG at key load (qjit.cu). Host-only check: the
    text for a few AIRs (base-field only, LogUp / extension field, preprocessed + cached partitions, a multi-kernel
    split) assembles with ptxas for sm_100a."""
    import ctypes as C
    import shutil
    import subprocess
    import numpy as np
    from openvm_b200 import _lib, airs as A
    ptxas = shutil.which("ptxas") or "/usr/local/cuda/bin/ptxas"
    if not os.path.exists(ptxas):
        pytest.skip("no ptxas")
    lib = _lib.load()
    a_parts = A.lookup_chain_air_parts(6, 64, seed=2)
    a_parts.prep_commit = np.arange(8, dtype=np.uint32)   # host-only parse: the commitment is not checked
    cases = [[A.fibonacci_air()], [A.segment_air(8)], [a_parts, A.multi_receiver_air([2, 3], 5)], [A.segment_air(64)]]
    os.environ["B2S_QUOTIENT_JIT_LINES"] = "3000"
    try:
        for air_list in cases:
            key = A.key_blob(air_list, log_up_pow_bits=1)
            for i in range(len(air_list)):
                out = C.c_void_p()
                assert lib.b2s_debug_quotient_ptx(key.ctypes.data, key.size, 1, i, C.byref(out)) == 0
                text = C.string_at(out.value).decode()
                lib.b2s_free(out)
                mods = [m for m in text.split("// ---- next module\n") if m.strip()]
                assert mods and all(".entry qk(" in m for m in mods)
                for j, m in enumerate(mods):
                    f = tmp_path / f"q{i}_{j}.ptx"
                    f.write_text(m)
                    r = subprocess.run([ptxas, "-arch=sm_100a", "-O1", str(f), "-o", str(f) + ".cubin"], capture_output=True, text=True)
                    assert r.returncode == 0, r.stderr[:2000]
        assert len(mods) > 1   # the last case is split into several kernels
    finally:
        del os.environ["B2S_QUOTIENT_JIT_LINES"]


def test_rust_ffi_bindings_are_current_and_complete():
    """integration/b2stark-engine/src/ffi.rs (the binding a maintainer compiles into the reference) is the
    output of tools/gen_ffi_rs.py on the current header and declares every function the header declares."""
    import re
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_ffi_rs as g
    text = open(g.OUT).read()
    assert text == g.render()
    header = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "b2stark.h")).read(), flags=re.S)
    declared = set(re.findall(r"\b(b2s_\w+)\s*\(", header))
    bound = set(re.findall(r"pub fn (b2s_\w+)\(", text))
    assert declared == bound and len(bound) >= 50
    # every wrapper call in lib.rs goes to a bound symbol
    lib_rs = open(os.path.join(ROOT, "integration", "b2stark-engine", "src", "lib.rs")).read()
    assert set(re.findall(r"ffi::(b2s_\w+)", lib_rs)) <= bound
