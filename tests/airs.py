"""Test-side alias of openvm_b200.airs (key-blob writer and sample AIRs)."""
from openvm_b200.airs import *  # noqa: F401,F403
from openvm_b200.airs import Air, Dag, key_blob, fibonacci_air, fibonacci_trace, chain_air, chain_trace  # noqa: F401
