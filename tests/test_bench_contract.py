"""bench.py on a CPU-only box: the reference arm prints one JSON line with the contract's keys, and the bench's own arithmetic
(algorithmic bytes, permutation counts, M_Q of a DAG) matches SURVEY.md 8(d)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_reference_arm_line_on_cpu():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--log-height", "12", "--width", "16",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "lde_commit_throughput" and d["unit"] == "GB/s" and d["value"] > 0
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert d["config"]["same_config"] is True and "2^12" in d["config"]["workload"]
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"], env=env,
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_bench_arithmetic():
    import bench
    from openvm_b200 import airs
    # A_commit = (1 + 2^b) N W 4 + 2^(b+1) N 32 ; P_MMCS = 2^b N ceil(W/8) + 2^b N - 1   
    assert bench.algorithmic_bytes(24, 256, 1) == 3 * (1 << 24) * 256 * 4 + 4 * (1 << 24) * 32 == 53687091200
    assert bench.permutations(24, 256, 1) == (1 << 25) * 32 + (1 << 25) - 1
    assert bench.permutations(10, 19, 2) == (1 << 12) * 3 + (1 << 12) - 1
    # M_Q weights: Fibonacci AIR = five selector x difference products (F x F) and five constraints
    fib = airs.fibonacci_air()
    assert bench.dag_weights(fib) == 5 * 1 + 16 * 5
    # an AIR with LogUp has EF x F and EF x EF products
    a = airs.dummy_interaction_air(2, True)
    w = bench.dag_weights(a)
    assert w > 16 * len(a.dag.constraints) and (w - 16 * len(a.dag.constraints)) >= 16
    kinds = [k for k, *_ in a.dag.nodes]
    assert airs.N_PERM in kinds and airs.N_CHALLENGE in kinds
