"""Population-level sharding of fitness evaluations over the GPUs of one node.

Mirrors the reference's only parallelism strategy: data-parallel evaluation of the GP population
over MPI ranks (/root/reference/evostencils/optimization/program.py:285-310 `allgather`/`gather`,
:478-502 initial population, :534-535 re-evaluation striding `i % nprocs == rank`).  Here one
process per GPU (torchrun) owns one `B200ProgramGenerator`; individual `i` of the de-duplicated
population is evaluated by rank `i % world`; the three doubles per individual
`(time_ms, convergence_factor, iterations)` are combined on every rank.  Evaluations are
independent, so there is no data-path collective -- only this result exchange.

`torch.distributed` is plumbing: NCCL on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

import contextlib
from typing import Callable, List, Optional, Sequence, Tuple

Result = Tuple[float, float, float]


def shard_indices(n_items: int, rank: int, world: int) -> List[int]:
    """Indices owned by `rank`: the reference's striding (program.py:534-535)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    return list(range(rank, n_items, world))


def deduplicate(strings: Sequence[str]):
    """Unique individuals in first-seen order + the map back.  The reference's cache key is
    `str(individual)` (program.py:191-204): equal strings are evaluated once."""
    unique: List[str] = []
    index = {}
    back = []
    for s in strings:
        if s not in index:
            index[s] = len(unique)
            unique.append(s)
        back.append(index[s])
    return unique, back


def _world(group=None):
    try:
        import torch.distributed as dist
    except Exception:       # pragma: no cover
        return 0, 1, None
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group), dist
    return 0, 1, None


class WorkQueue:
    """Self-scheduling of one batch of jobs over the ranks: a rank takes the next job index from an
    atomic counter kept in the process group's key-value store (`Store.add`, a TCP round trip of
    ~0.1 ms to rank 0 against tens of ms per evaluation).  The reference deals `i % nprocs`
    (program.py:534-535), which leaves ranks idle when evaluation times differ (individuals that
    converge in 15 cycles next to ones that run all 100); the set of individuals, the result per
    individual and the exchange are unchanged, only *which* rank evaluates an individual.

    Every rank must create its queues in the same order (the n-th queue of every rank shares a key)."""

    _serial = 0

    def __init__(self, n_jobs: int, group=None, store=None):
        import threading
        rank, world, dist = _world(group)
        self.n_jobs = n_jobs
        self.world = world
        self._next = 0
        self._lock = threading.Lock()           # several evaluation threads of one rank may share a queue
        self._store = None
        if world > 1:
            if store is None:
                from torch.distributed import distributed_c10d
                store = distributed_c10d._get_default_store()
            WorkQueue._serial += 1
            self._store = store
            self._key = "evostencils_b200/workqueue/%d" % WorkQueue._serial

    def __iter__(self):
        return self

    def __next__(self) -> int:
        with self._lock:
            if self._store is None:
                k = self._next
                self._next += 1
            else:
                k = self._store.add(self._key, 1) - 1
        if k >= self.n_jobs:
            raise StopIteration
        return k


def gather_results(local: dict, n_items: int, group=None, device=None, error: Optional[BaseException] = None) -> List[Result]:
    """Combine `{index: (t, rho, n)}` of every rank into the full list on every rank.

    Each slot is written by exactly one rank and is 0.0 elsewhere, so the SUM all-reduce is exact
    (x + 0.0 == x for every finite x and for +inf) and independent of the reduction order.
    `error`: the exception that stopped this rank's evaluations, if any.  The rank still takes part
    in the exchange (an extra row counts the failed ranks), so the others are not left waiting in
    the collective; afterwards every rank raises."""
    rank, world, dist = _world(group)
    if world == 1:
        if error is not None:
            raise error
        missing = [i for i in range(n_items) if i not in local]
        if missing:
            raise ValueError(f"results missing for individuals {missing[:5]}")
        return [tuple(float(v) for v in local[i]) for i in range(n_items)]
    import torch
    if device is None:
        device = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    buf = torch.zeros((n_items + 1, 4), dtype=torch.float64)
    for i, r in local.items():
        buf[i, 0], buf[i, 1], buf[i, 2], buf[i, 3] = float(r[0]), float(r[1]), float(r[2]), 1.0
    buf[n_items, 0] = 0.0 if error is None else 1.0
    buf = buf.to(device)
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    buf = buf.cpu()
    failed = int(buf[n_items, 0].item())
    if error is not None:
        raise error
    if failed:
        raise RuntimeError(f"the evaluation failed on {failed} other rank(s)")
    buf = buf[:n_items]
    counts = buf[:, 3]
    if not bool((counts == 1.0).all()):
        bad = [i for i in range(n_items) if counts[i] != 1.0]
        raise RuntimeError(f"individuals evaluated {counts[bad[0]].item():.0f} times (expected once): {bad[:5]}")
    return [(buf[i, 0].item(), buf[i, 1].item(), buf[i, 2].item()) for i in range(n_items)]


def run_concurrently(evaluators: Sequence[Callable[[str], Result]], jobs, unique: Sequence[str]) -> dict:
    """Evaluate `unique[j]` for every j the iterator `jobs` yields, with one host thread per evaluator
    (= per engine workspace and CUDA stream of this rank).  The engine calls release the GIL, so the
    solves of the threads overlap on the GPU: the latency-bound coarse levels of one individual run
    under the bandwidth-bound fine levels of another (`north_star`: one individual per stream).
    Results per individual do not depend on the overlap; the measured solve *time* of an individual
    does (it shares the GPU), like a reference run that shares a node's memory bandwidth with other
    MPI ranks.  -> {j: result}"""
    import threading
    local: dict = {}
    errors: list = []
    lock = threading.Lock()
    it = iter(jobs)

    def take():
        with lock:
            return next(it, None)

    def work(evaluate):
        try:
            while True:
                j = take()
                if j is None:
                    return
                r = evaluate(unique[j])
                with lock:
                    local[j] = r
        except BaseException as exc:            # surfaced on the calling thread
            errors.append(exc)

    if len(evaluators) == 1:
        work(evaluators[0])
    else:
        threads = [threading.Thread(target=work, args=(e,), daemon=True) for e in evaluators]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    if errors:
        errors[0].partial_results = local
        raise errors[0]
    return local


def evaluate_population(evaluate, strings: Sequence[str], group=None, device=None,
                        costs: Optional[Sequence[float]] = None, schedule: str = "static") -> List[Result]:
    """Evaluate every individual once across the ranks of `group`; every rank gets all results.
    `evaluate` may be a list of callables, one per engine workspace of this rank: they run in
    parallel host threads and share the rank's part of the population (`run_concurrently`).

    `evaluate(grammar_string) -> (time_ms, rho, n)` runs on this rank's GPU.  With `costs`
    (e.g. op-list byte counts) the unique individuals are dealt longest-first so that the ranks
    finish together; without, plain `i % world` striding.  `schedule="dynamic"`
    replaces the striding by a `WorkQueue`: ranks take the next individual of the (cost-ordered)
    list when they finish one."""
    rank, world, _ = _world(group)
    unique, back = deduplicate(strings)
    order = list(range(len(unique)))
    if costs is not None:
        if len(costs) != len(strings):
            raise ValueError("one cost per individual expected")
        first = {}
        for i, b in enumerate(back):
            first.setdefault(b, costs[i])
        order.sort(key=lambda u: (-first[u], u))
    evaluators = list(evaluate) if isinstance(evaluate, (list, tuple)) else [evaluate]
    if schedule == "dynamic":
        jobs = (order[k] for k in WorkQueue(len(order), group))
    elif schedule == "static":
        jobs = [order[k] for k in shard_indices(len(order), rank, world)]
    else:
        raise ValueError("schedule must be 'static' or 'dynamic'")
    error = None
    try:
        local = run_concurrently(evaluators, jobs, unique)
    except Exception as exc:            # still join the exchange, then raise on every rank
        error, local = exc, getattr(exc, "partial_results", {})
    results = gather_results(local, len(unique), group, device, error)
    return [results[b] for b in back]


def make_evaluator(generator, grammar, min_level: int, max_level: int, evaluation_samples: int = 1,
                   infinity: float = 1e100) -> Callable[[str], Result]:
    """`evaluate` for `evaluate_population`: grammar string -> Cycle graph -> generate_and_evaluate
    (the path of program.py:904-929 `generate_and_evaluate_program_from_grammar_representation`)."""
    def evaluate(string: str) -> Result:
        try:
            expression, _ = grammar.compile(string)
        except Exception:
            return (infinity, infinity, infinity)
        return tuple(generator.generate_and_evaluate(expression, None, min_level, max_level, "",
                                                     infinity=infinity, evaluation_samples=evaluation_samples))
    return evaluate


class SoloGate:
    """Readers-writer gate shared by the generators of one rank: ordinary solves hold it `shared`
    (they overlap on the GPU), the timing replay of an individual holds it `exclusive` (nothing else
    of this process runs on the GPU meanwhile)."""

    def __init__(self):
        import threading
        self._cond = threading.Condition()
        self._readers = 0
        self._writer = False
        self._writers_waiting = 0

    @contextlib.contextmanager
    def shared(self):
        with self._cond:
            while self._writer or self._writers_waiting:
                self._cond.wait()
            self._readers += 1
        try:
            yield
        finally:
            with self._cond:
                self._readers -= 1
                self._cond.notify_all()

    @contextlib.contextmanager
    def exclusive(self):
        with self._cond:
            self._writers_waiting += 1
            while self._writer or self._readers:
                self._cond.wait()
            self._writers_waiting -= 1
            self._writer = True
        try:
            yield
        finally:
            with self._cond:
                self._writer = False
                self._cond.notify_all()


def node_count(individual) -> int:
    """Number of nodes of a derivation tree: `len(individual)` of a deap PrimitiveTree; for a grammar
    string one per call, name and literal (a block-shape tuple is one terminal, multigrid.py:388-407)."""
    if not isinstance(individual, str):
        return len(individual)
    import ast

    def count(node):
        if isinstance(node, ast.Call):
            return 1 + sum(count(a) for a in node.args)
        return 1
    return count(ast.parse(individual, mode="eval").body)


def fitness_from_result(result: Result, objectives: str = "multi", infinity: float = 1e100):
    """(time_ms, rho, n) -> the fitness tuple of the reference's wrappers: `multi` =
    evaluate_multiple_objectives (program.py:419-453): (rho, t / n), or (rho, infinity) when the solver
    did not converge; `single` = evaluate_single_objective (:386-417): (t,), or (sqrt(rho) sqrt(n),)."""
    t, rho, n = result
    if objectives == "single":
        if n >= infinity:
            return (rho ** 0.5 * n ** 0.5,)
        return (t,)
    if objectives != "multi":
        raise ValueError("objectives must be 'multi' or 'single'")
    if n >= infinity:
        return (rho, infinity)
    return (rho, t / n)


MAX_INDIVIDUAL_SIZE = 150            # program.py:390, 423: larger trees are not evaluated


def batched_map(generator, grammar, min_level, max_level, evaluation_samples=1, group=None, infinity=1e100,
                schedule="static", objectives=None, solo_timing=True):
    """A `toolbox.map` replacement (the reference's extension point, program.py:491, 577):
    `toolbox.register("map", batched_map(...))` makes one generation's evaluations run sharded.
    Individuals (grammar strings / PrimitiveTrees) are evaluated through `generate_and_evaluate` and
    mapped to the fitness tuple of the wrapper that `toolbox.map` was called with: the function's name
    (`evaluate_single_objective` / `evaluate_multiple_objectives`, also through functools.partial) or
    the explicit `objectives` = "single" | "multi" selects the mapping (program.py:386-453); trees of
    more than 150 nodes get the infinity tuple without being evaluated (:390, :423).

    `solo_timing`: with several generators per rank the solves overlap on the GPU for throughput, which
    makes the measured time of an individual depend on its neighbour; the time objective is therefore
    taken from a replay of the already captured cycle graph while the rank's other generators pause
    (`SoloGate`, `B200ProgramGenerator.solo_timing`): time = replayed ms per cycle x iterations.
    False keeps the overlapped timings.  One generator per rank: solves are solo anyway."""
    generators = list(generator) if isinstance(generator, (list, tuple)) else [generator]
    if solo_timing and len(generators) > 1:
        gate = SoloGate()
        for g in generators:
            g.solo_timing = gate
    # several generators (one workspace and stream each): their evaluations overlap on the GPU
    evaluators = [make_evaluator(g, grammar, min_level, max_level, evaluation_samples, infinity) for g in generators]

    def _objectives(function):
        if objectives is not None:
            return objectives
        f = function
        while hasattr(f, "func"):          # functools.partial
            f = f.func
        return "single" if "single" in getattr(f, "__name__", "") else "multi"

    def _map(function, individuals):
        individuals = list(individuals)
        kind = _objectives(function)
        too_large = (infinity,) if kind == "single" else (infinity, infinity)
        keep = [i for i, ind in enumerate(individuals) if node_count(ind) <= MAX_INDIVIDUAL_SIZE]
        strings = [str(individuals[i]) for i in keep]
        out = [too_large] * len(individuals)
        for i, result in zip(keep, evaluate_population(evaluators, strings, group, schedule=schedule)):
            out[i] = fitness_from_result(result, kind, infinity)
        return out
    return _map
