"""Root configuration (BabyBearPoseidon2RootConfig): the chain-AIR segment shape proved with the BN254 hash, next to the same
proof under the inner configuration; plus the raw rate of the device Poseidon2-BN254 permutation.
python tools/root_shape.py [log_h] [width]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from openvm_b200 import B2Engine, FriParameters, AirProvingContext, airs as A

log_h = int(sys.argv[1]) if len(sys.argv) > 1 else 16
width = int(sys.argv[2]) if len(sys.argv) > 2 else 64
air = A.segment_air(width)
key = A.build_key([air], None, log_up_pow_bits=16)
parts, pv = A.chain_trace(air, 1 << log_h, seed=4)
for kind, name in ((0, "inner (Poseidon2-BabyBear-16)"), (1, "root  (Poseidon2-BN254-3)")):
    eng = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(2))
    eng.set_hash(kind)
    dev = eng.device()
    pk = dev.transport_pk_to_device(key)
    ctx = [AirProvingContext(0, dev.transport_matrix_to_device(parts[-1]), pv)]
    for it in range(3):
        eng.reset_metrics()
        t = time.perf_counter()
        proof = eng.prove(pk, ctx)
        wall = (time.perf_counter() - t) * 1e3
        m = eng.metrics()
    print(f"{name}: 2^{log_h} x {width}: proof wall {wall:.1f} ms, commit {m['commit_total_ms']:.1f} (lde {m['lde_ms']:.1f}, leaf {m['leaf_hash_ms']:.1f}, "
          f"compress {m['compress_ms']:.1f}), quotient+logup {(m['quotient_poly_compute_time_ms'] + m['generate_perm_trace_time_ms']):.1f}, opening {m['pcs_opening_time_ms']:.1f}, "
          f"{m['poseidon2_permutations']} permutations, {len(proof)} bytes", flush=True)
    eng.verify(pk, proof)
    if kind == 1:
        n = 1 << 20
        st = np.zeros((n, 24), dtype=np.uint32)
        st[:, 0] = np.arange(n)
        dev.poseidon2_bn254_permute(st[:1024])
        import ctypes
        t = time.perf_counter()
        dev.poseidon2_bn254_permute(st)
        dt = time.perf_counter() - t
        print(f"device Poseidon2-BN254: 2^20 permutations incl. 192 MB of copies in {dt * 1e3:.1f} ms")
    eng.close()
