import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openvm_b200 import B2Engine, FriParameters
eng = B2Engine.new(FriParameters(log_blowup=1))
n = 148 * 2048
for mask in (0, 1, 4, 5, 8, 9, 13, 15):
    os.environ["B2S_P2_MASK"] = str(mask)
    eng.bench_poseidon2(n, 16)
    ms = min(eng.bench_poseidon2(n, 64) for _ in range(3))
    print(f"mask {mask:2d}: {n * 64 / ms / 1e6:.3f} G perm/s", flush=True)
eng.close()
