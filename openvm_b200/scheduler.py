"""Host-side partitioning of independent proving jobs across GPUs.

Continuation segments and aggregation-tree nodes are independent STARK proofs (the reference proves them in a
sequential loop on one engine, crates/vm/src/arch/vm.rs:979-1007, and documents the distributed model in
docs/vocs/docs/pages/specs/architecture/distributed-proving.mdx). One process per GPU, one engine per process; the
only cross-rank traffic is a barrier, the max-over-ranks time and the gather of proof sizes -- no data-path collective.
"""
import torch.distributed as dist


def partition(n_jobs, world_size, rank):
    """Job indices of `rank`: round-robin, so layer k of an arity-3 aggregation tree (ceil(S / 3^k) nodes) stays balanced."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    return list(range(rank, n_jobs, world_size))


def aggregation_layers(n_segments, num_children_leaf=1, num_children_internal=3):
    """Number of independent jobs per layer: leaves (1:1 with segments by default, crates/sdk/src/config/mod.rs:26), then
    internal layers of arity 3 (:27) until one proof is left."""
    layers = [-(-n_segments // num_children_leaf)]
    while layers[-1] > 1:
        layers.append(-(-layers[-1] // num_children_internal))
    return layers


def max_over_ranks(value, device=None):
    """Max of a python float over all ranks (time-like quantities are reported as the slowest rank's)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_counts(value):
    """All ranks' integer `value` (e.g. segments proven), on every rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [int(value)]
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, int(value))
    return out


class ProverPool:
    """K engines on one GPU, one host thread each, fed from one FIFO of proving jobs with dependencies.

    Why: a proof alternates host work (transcript, round trips) with kernels too small to fill a B200, so several proofs in
    flight raise segments/s (one engine 50/s, eight 81/s for the RV32IM-shaped segment) -- the reference keeps
    one engine per VM kind on a GPU (crates/sdk/src/prover/stark.rs:62-75) and proves segments one after the other
    (crates/vm/src/arch/vm.rs:979-1007). Jobs are the nodes of the continuation / aggregation DAG (segment i, leaf i after
    segment i, an internal node after its <= 3 children: SURVEY.md 8e); a job becomes ready when its dependencies are done
    and receives their results.

        pool = ProverPool(lambda k: make_worker(k), engines=8)
        a = pool.submit(lambda worker: worker.prove_segment(0))
        b = pool.submit(lambda worker, proof: worker.prove_leaf(proof), deps=[a])
        proofs = pool.results([a, b]); pool.close()

    `make_worker(k)` runs on worker thread k and returns whatever the jobs need (typically an object holding a B2Engine, its
    device keys and resident traces); the pool never touches CUDA itself, so it is tested on CPU with plain callables."""

    def __init__(self, make_worker, engines=8):
        import queue
        import threading
        if engines < 1:
            raise ValueError("engines must be >= 1")
        self._lock = threading.Condition()
        self._ready = queue.Queue()
        self._jobs = {}          # id -> [fn, deps, state, result]  state: 0 waiting, 1 queued/running, 2 done, 3 failed
        self._waiters = {}       # dep id -> [job ids]
        self._next = 0
        self._closed = False
        self._threads = [threading.Thread(target=self._run, args=(make_worker, k), daemon=True) for k in range(engines)]
        self._startup_error = None
        for t in self._threads:
            t.start()

    def submit(self, fn, deps=()):
        """Queue `fn(worker, *results_of_deps)`; returns the job id. A job whose dependency failed fails with it."""
        with self._lock:
            if self._closed:
                raise RuntimeError("pool is closed")
            jid = self._next
            self._next += 1
            deps = list(deps)
            for d in deps:
                if d not in self._jobs:
                    raise ValueError("unknown dependency %r" % (d,))
            self._jobs[jid] = [fn, deps, 0, None]
            pending = [d for d in deps if self._jobs[d][2] < 2]
            for d in pending:
                self._waiters.setdefault(d, []).append(jid)
            if not pending:
                self._release(jid)
            return jid

    def _release(self, jid):       # lock held: every dependency is finished
        job = self._jobs[jid]
        failed = [d for d in job[1] if self._jobs[d][2] == 3]
        if failed:
            self._finish(jid, 3, self._jobs[failed[0]][3])
        else:
            job[2] = 1
            self._ready.put(jid)

    def _finish(self, jid, state, result):   # lock held
        job = self._jobs[jid]
        job[2], job[3] = state, result
        for w in self._waiters.pop(jid, []):
            if all(self._jobs[d][2] >= 2 for d in self._jobs[w][1]):
                self._release(w)
        self._lock.notify_all()

    def _run(self, make_worker, k):
        try:
            worker = make_worker(k)
        except BaseException as ex:            # noqa: BLE001 - reported to whoever waits
            with self._lock:
                self._startup_error = ex
                self._lock.notify_all()
            worker = None
        while True:
            jid = self._ready.get()
            if jid is None:
                return
            with self._lock:
                fn, deps = self._jobs[jid][0], self._jobs[jid][1]
                args = [self._jobs[d][3] for d in deps]
            try:
                if worker is None:
                    raise RuntimeError("worker %d failed to start: %r" % (k, self._startup_error))
                out, state = fn(worker, *args), 2
            except BaseException as ex:        # noqa: BLE001
                out, state = ex, 3
            with self._lock:
                self._finish(jid, state, out)

    def result(self, jid, timeout=None):
        """Blocks until job `jid` is done; returns its result or raises its exception."""
        with self._lock:
            if not self._lock.wait_for(lambda: self._jobs[jid][2] >= 2, timeout):
                raise TimeoutError("job %d still running" % jid)
            state, out = self._jobs[jid][2], self._jobs[jid][3]
        if state == 3:
            raise out
        return out

    def results(self, jids, timeout=None):
        return [self.result(j, timeout) for j in jids]

    def close(self):
        with self._lock:
            self._closed = True
        for _ in self._threads:
            self._ready.put(None)
        for t in self._threads:
            t.join()


def submit_aggregation_tree(pool, n_segments, prove_segment, prove_leaf, prove_internal, num_children_internal=3):
    """Queues the whole continuation DAG on `pool`: segment i; leaf i after segment i (1:1,
    crates/sdk/src/config/mod.rs:26); internal nodes of arity 3 (:27) layer by layer until one proof is left. The callables
    get (worker, index, *child_proofs). Returns (root job id, all job ids by layer)."""
    segs = [pool.submit(lambda w, i=i: prove_segment(w, i)) for i in range(n_segments)]
    layer = [pool.submit(lambda w, p, i=i: prove_leaf(w, i, p), deps=[s]) for i, s in enumerate(segs)]
    layers = [segs, layer]
    while len(layer) > 1:
        nxt = []
        for i in range(0, len(layer), num_children_internal):
            kids = layer[i:i + num_children_internal]
            nxt.append(pool.submit(lambda w, *ps, i=i // num_children_internal: prove_internal(w, i, *ps), deps=kids))
        layer = nxt
        layers.append(layer)
    return layer[0], layers


def aggregation_dag(n_segments, num_children_internal=3):
    """The continuation DAG as (kinds, deps): job ids in FIFO order -- segments 0..S-1, leaf S+i after segment i (1:1,
    crates/sdk/src/config/mod.rs:26), then internal layers of arity 3 (:27; crates/sdk/src/prover/agg.rs:124-202) until one
    proof is left. kinds[j] = ("segment" | "leaf" | "internal", index within its layer, layer number)."""
    kinds = [("segment", i, 0) for i in range(n_segments)]
    deps = [[] for _ in range(n_segments)]
    layer = []
    for i in range(n_segments):
        layer.append(len(kinds))
        kinds.append(("leaf", i, 1))
        deps.append([i])
    depth = 2
    while len(layer) > 1:
        nxt = []
        for i in range(0, len(layer), num_children_internal):
            nxt.append(len(kinds))
            kinds.append(("internal", i // num_children_internal, depth))
            deps.append(layer[i:i + num_children_internal])
        layer = nxt
        depth += 1
    return kinds, deps


def critical_path(deps, cost):
    """Length of the longest dependency chain when job j costs cost[j] (the lower bound of the DAG's wall time)."""
    done = [0.0] * len(deps)
    for j, d in enumerate(deps):
        done[j] = cost[j] + max((done[k] for k in d), default=0.0)
    return max(done, default=0.0)


class DagFifo:
    """ONE FIFO of DAG jobs shared by every engine of every rank ("job -> first free engine"), over a
    torch.distributed key-value store (the TCPStore the process group already has; a HashStore in one process). No
    data-path collective: a worker claims the next job id with an atomic counter, waits until the results of its
    dependencies are in the store, runs it, and publishes its result (proof bytes: what the parent's verifier program is fed,
    crates/sdk/src/prover/agg.rs:307-321). Job ids are topologically ordered (aggregation_dag), so the unfinished job with
    the smallest id never waits on an unclaimed one: no deadlock, and an engine only idles on a dependency at the tail of
    the DAG, where the critical path is the limit anyway.

        fifo = DagFifo(store, deps, tag="run0")
        log = fifo.work(lambda jid, child_results: prove(jid, child_results))   # on every engine thread of every rank
    """

    def __init__(self, store, deps, tag="dag", poll_s=0.0005):
        self.store, self.deps, self.tag, self.poll_s = store, deps, tag, poll_s

    def _key(self, what, j=None):
        return f"{self.tag}/{what}" if j is None else f"{self.tag}/{what}/{j}"

    def work(self, fn, clock=None):
        """Claims and runs jobs until the FIFO is empty. fn(jid, [results of deps]) -> bytes. Returns this worker's log:
        [(jid, t_claimed, t_ready, t_done)] on `clock` (default time.perf_counter)."""
        import time
        clock = clock or time.perf_counter
        log = []
        n = len(self.deps)
        while True:
            jid = self.store.add(self._key("next"), 1) - 1
            if jid >= n:
                return log
            t0 = clock()
            keys = [self._key("done", d) for d in self.deps[jid]]
            # poll instead of store.wait(): a blocking wait holds the store client of this PROCESS, and the engine thread
            # that will publish the dependency may live in the same process
            while keys and not self.store.check(keys):
                time.sleep(self.poll_s)
            args = [bytes(self.store.get(k)) for k in keys]
            t1 = clock()
            out = fn(jid, args)
            self.store.set(self._key("done", jid), out if isinstance(out, (bytes, bytearray)) else bytes(out))
            log.append((jid, t0, t1, clock()))

    def result(self, jid):
        return bytes(self.store.get(self._key("done", jid)))
