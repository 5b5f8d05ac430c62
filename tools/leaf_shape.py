"""Config 5: shape replay of a leaf-aggregation proof -- 15 AIRs, heights up to 2^22, ~446 M main cells,
quotient degree up to 4, FRI log_blowup 2, LogUp on every AIR. Authored chain AIRs with satisfying traces; proves on the
device, checks with b2s_verify, prints stage times.  python tools/leaf_shape.py [scale_log] [proofs]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from openvm_b200 import B2Engine, FriParameters, AirProvingContext, airs as A

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 0
proofs = int(sys.argv[2]) if len(sys.argv) > 2 else 3
b = 2
shapes = A.leaf_shapes(scale)
t = time.perf_counter()
air_objs = [A.segment_air(w, degree=dg) for _, w, dg in shapes]
key = A.build_key(air_objs, None, log_up_pow_bits=16)
print(f"{len(air_objs)} AIRs, key {key.nbytes / 1e6:.1f} MB, built in {time.perf_counter() - t:.1f} s", flush=True)
eng = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(b))
dev = eng.device()
t = time.perf_counter()
pk = dev.transport_pk_to_device(key)
print(f"key load (with code generation): {time.perf_counter() - t:.2f} s", flush=True)
ctx, cells, perm_cells = [], 0, 0
t = time.perf_counter()
for i, (lh, w, dg) in enumerate(shapes):
    parts, pv = A.chain_trace(air_objs[i], 1 << lh, seed=7 + i)
    ctx.append(AirProvingContext(i, dev.transport_matrix_to_device(parts[-1]), pv))
    cells += w << lh
    perm_cells += (4 * (len(air_objs[i].chunks) + 1)) << lh
    del parts
print(f"traces: {cells / 1e6:.1f} M main cells, {perm_cells / 1e6:.1f} M after-challenge cells, generated and uploaded in "
      f"{time.perf_counter() - t:.1f} s", flush=True)
for it in range(proofs):
    eng.reset_metrics()
    t = time.perf_counter()
    proof = eng.prove(pk, ctx)
    wall = (time.perf_counter() - t) * 1e3
    m = eng.metrics()
    print(f"proof {it}: wall {wall:.1f} ms, commit {m['commit_total_ms']:.1f} (lde {m['lde_ms']:.1f}, leaf {m['leaf_hash_ms']:.1f}), "
          f"quotient+logup {(m['quotient_poly_compute_time_ms'] + m['generate_perm_trace_time_ms']):.1f}, opening {m['pcs_opening_time_ms']:.1f}, {len(proof)} bytes, "
          f"device peak {eng.device_memory()[1] / 2**30:.1f} GiB", flush=True)
t = time.perf_counter()
eng.verify(pk, proof)
print(f"b2s_verify: accepted in {(time.perf_counter() - t) * 1e3:.1f} ms")
eng.close()
