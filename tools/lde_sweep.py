"""LDE timing sweep (GPU box): python tools/lde_sweep.py [log_h ...]  -- prints lde ms per config and group size."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openvm_b200 import B2Engine, FriParameters

def run(log_h, width, b, reps=5):
    eng = B2Engine.new(FriParameters(log_blowup=b))
    dev = eng.device()
    m = dev.alloc_matrix(1 << log_h, width)
    dev.fill_synthetic(m, 1)
    for _ in range(2):
        dev.coset_lde(m, b).free()
    eng.reset_metrics()
    for _ in range(reps):
        dev.coset_lde(m, b).free()
    met = eng.metrics()
    ms = met["lde_ms"] / reps
    n = (1 << log_h) * width
    bf = n * log_h / 2 * (1 + (1 << b))
    print(f"log_h={log_h} w={width} b={b} group={os.environ.get('B2S_LDE_GROUP_ELEMS','default')} plan={os.environ.get('B2S_NTT_PLAN','default')}: {ms:.3f} ms, "
          f"{bf / ms / 1e6:.1f} G butterflies/s, launches/rep {met['lde_launches'] / reps:.0f}", flush=True)
    m.free(); eng.close()

if __name__ == "__main__":
    hs = [int(x) for x in sys.argv[1:]] or [20]
    for h in hs:
        run(h, 256 if h <= 22 else 64, 1)
