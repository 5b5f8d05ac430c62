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
