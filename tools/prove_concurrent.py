"""Segments/s with K engines proving concurrently on one GPU (one host thread each): python tools/prove_concurrent.py [K] [n]"""
import os, sys, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openvm_b200 import B2Engine, FriParameters, AirProvingContext, airs as A

K = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n = int(sys.argv[2]) if len(sys.argv) > 2 else 20
shapes = A.segment_shapes(0)
air_objs = [A.segment_air(w, perm_cols=pc) for _, w, pc in shapes]
key = A.build_key(air_objs, None, log_up_pow_bits=16)
traces = [A.chain_trace(air_objs[i], 1 << lh, seed=i) for i, (lh, w, _pc) in enumerate(shapes)]
engines = []
for k in range(K):
    eng = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(1))
    dev = eng.device()
    pk = dev.transport_pk_to_device(key)
    ctx = [AirProvingContext(i, dev.transport_matrix_to_device(p[-1]), pv) for i, (p, pv) in enumerate(traces)]
    eng.prove(pk, ctx)
    engines.append((eng, pk, ctx))

def work(e):
    eng, pk, ctx = e
    for _ in range(n):
        eng.prove(pk, ctx)

for kk in range(1, K + 1):
    ths = [threading.Thread(target=work, args=(engines[i],)) for i in range(kk)]
    t = time.perf_counter()
    for th in ths: th.start()
    for th in ths: th.join()
    dt = time.perf_counter() - t
    print(f"{kk} engine(s): {kk * n / dt:.1f} segments/s ({dt / n * 1e3:.2f} ms per round of {kk})", flush=True)
for eng, _, _ in engines:
    eng.close()
