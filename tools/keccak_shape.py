"""Config 4: wide, quotient-heavy AIR with the KeccakVmAir shape -- 2^17 rows x 3164 main columns,
~4.6 k constraints of degree <= 3 (quotient degree 2) and LogUp traffic worth ~320 extension columns. Authored chain
AIR, satisfying trace; proves on the device, checks with b2s_verify, prints stage times.  python tools/keccak_shape.py [log_h] [width]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from openvm_b200 import B2Engine, FriParameters, AirProvingContext, airs as A

log_h = int(sys.argv[1]) if len(sys.argv) > 1 else 17
width = int(sys.argv[2]) if len(sys.argv) > 2 else 3164
t = time.perf_counter()
air = A.segment_air(width)
key = A.build_key([air], None, log_up_pow_bits=16)
print(f"air: {len(air.dag.nodes)} nodes, {len(air.dag.constraints)} constraints, {len(air.interactions)} interactions, "
      f"{len(air.chunks) + 1} EF after-challenge columns, key {key.nbytes / 1e6:.1f} MB, built in {time.perf_counter() - t:.1f} s", flush=True)
t = time.perf_counter()
parts, pv = A.chain_trace(air, 1 << log_h, seed=4)
print(f"trace {parts[-1].shape} in {time.perf_counter() - t:.1f} s", flush=True)
eng = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(1))
dev = eng.device()
pk = dev.transport_pk_to_device(key)
ctx = [AirProvingContext(0, dev.transport_matrix_to_device(parts[-1]), pv)]
for it in range(3):
    eng.reset_metrics()
    t = time.perf_counter()
    proof = eng.prove(pk, ctx)
    wall = (time.perf_counter() - t) * 1e3
    m = eng.metrics()
    print(f"proof {it}: wall {wall:.1f} ms, commit {m['commit_total_ms']:.1f} (lde {m['lde_ms']:.1f}, leaf {m['leaf_hash_ms']:.1f}), "
          f"quotient+logup {(m['quotient_poly_compute_time_ms'] + m['generate_perm_trace_time_ms']):.1f}, opening {m['pcs_opening_time_ms']:.1f}, {len(proof)} bytes", flush=True)
eng.verify(pk, proof)
print("b2s_verify: accepted")
eng.close()
