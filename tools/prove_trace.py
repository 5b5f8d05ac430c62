"""Host/device timeline of one shape-replay segment proof (B2S_PROVE_TRACE marks from prover.cu)."""
import os, sys, time
os.environ.setdefault("B2S_PROVE_TRACE", "2")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openvm_b200 import B2Engine, FriParameters, AirProvingContext, airs as A

eng = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(1))
dev = eng.device()
shapes = A.segment_shapes(int(sys.argv[1]) if len(sys.argv) > 1 else 0, legacy=len(sys.argv) > 2 and sys.argv[2] == "legacy")
air_objs = [A.segment_air(w, perm_cols=pc) for _, w, pc in shapes]
pk = dev.transport_pk_to_device(A.build_key(air_objs, None, log_up_pow_bits=16))
ctx = []
for i, (lh, w, _pc) in enumerate(shapes):
    m = dev.alloc_matrix(1 << lh, w)
    ctx.append(AirProvingContext(i, m, dev.fill_chain(m, 100 + i)))
for it in range(3):
    print(f"--- proof {it}", file=sys.stderr)
    t = time.perf_counter()
    proof = eng.prove(pk, ctx)
    print(f"wall {1e3 * (time.perf_counter() - t):.2f} ms, {len(proof)} bytes", file=sys.stderr)
eng.verify(pk, proof)
eng.close()
