"""Poseidon2 rates on the GPU box: register-resident permutation loop and the leaf-hash stage of a 2^20 x 256 commit."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openvm_b200 import B2Engine, FriParameters
eng = B2Engine.new(FriParameters(log_blowup=1))
n = 148 * 2048
eng.bench_poseidon2(n, 16)
ms = min(eng.bench_poseidon2(n, 64) for _ in range(3))
print(f"register loop: {n * 64 / ms / 1e6:.3f} G perm/s", flush=True)
dev = eng.device()
m = dev.alloc_matrix(1 << 20, 256)
dev.fill_synthetic(m, 1)
for _ in range(2):
    r, d = dev.commit([m]); d.free()
eng.reset_metrics()
reps = 5
for _ in range(reps):
    r, d = dev.commit([m]); d.free()
met = eng.metrics()
perms = (2 << 20) * 32 + (2 << 20) - 1
hash_ms = (met["leaf_hash_ms"] + met["compress_ms"]) / reps
print(f"commit 2^20 x 256: total {met['commit_total_ms'] / reps:.3f} ms, lde {met['lde_ms'] / reps:.3f}, leaf {met['leaf_hash_ms'] / reps:.3f}, "
      f"compress {met['compress_ms'] / reps:.3f}; {perms / hash_ms / 1e6:.3f} G perm/s", flush=True)
eng.close()
