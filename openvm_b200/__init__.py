"""openvm_b200 — host-side mirror of OpenVM's stark-backend prover interface over libb2stark.so.

The product is the C-ABI CUDA library (include/b2stark.h); this package is the thin host layer that mirrors
the reference's engine/device interface names (StarkFriEngine / ProverDevice, see engine.py). It never falls
back to a CPU path: importing works anywhere, but constructing an engine needs the built library and a B200.
"""
from .engine import (  # noqa: F401
    B2Engine,
    B2Device,
    DeviceMatrix,
    PcsData,
    ProvingKey,
    AirProvingContext,
    FriParameters,
    B2StarkError,
    VerificationError,
    library_path,
    verify_host,
)

__all__ = ["B2Engine", "B2Device", "DeviceMatrix", "PcsData", "ProvingKey", "AirProvingContext", "FriParameters", "B2StarkError", "VerificationError",
           "library_path", "verify_host"]
