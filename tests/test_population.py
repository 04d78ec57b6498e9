"""Population sharding: host logic on CPU, world_size 2 over gloo."""
import os
import socket
import sys

import pytest

from evostencils_b200 import population

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def fake_evaluate(string):
    """Deterministic stand-in for a GPU evaluation: depends on the string only."""
    h = sum(ord(c) * (i + 1) for i, c in enumerate(string))
    if h % 7 == 0:
        return (1e100, 1e100, 1e100)
    return (float(h % 1000) / 8.0, float(h % 97) / 97.0, float(h % 13 + 1))


STRINGS = ["ind_%d" % (i % 11) for i in range(29)]       # duplicates on purpose


def test_shard_indices_cover_everything_once():
    for world in (1, 2, 3, 8):
        seen = sorted(i for r in range(world) for i in population.shard_indices(29, r, world))
        assert seen == list(range(29))
    with pytest.raises(ValueError):
        population.shard_indices(4, 2, 2)


def test_deduplicate_is_first_seen_order():
    unique, back = population.deduplicate(STRINGS)
    assert unique == ["ind_%d" % i for i in range(11)]
    assert [unique[b] for b in back] == STRINGS


def test_single_process_matches_direct_evaluation():
    got = population.evaluate_population(fake_evaluate, STRINGS)
    assert got == [fake_evaluate(s) for s in STRINGS]
    costs = [float(len(s)) for s in STRINGS]
    assert population.evaluate_population(fake_evaluate, STRINGS, costs=costs) == got


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, with_costs, queue, schedule="static", evaluators=1):
    sys.path.insert(0, REPO)
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    calls = []

    def evaluate(s):
        calls.append(s)
        if schedule == "dynamic" and rank == 0:
            import time
            time.sleep(0.05)                  # a slow rank: the other one must take over its work
        return fake_evaluate(s)

    costs = [float((i * 7) % 5) for i in range(len(STRINGS))] if with_costs else None
    target = evaluate if evaluators == 1 else [evaluate] * evaluators        # several host threads per rank
    out = population.evaluate_population(target, STRINGS, costs=costs, schedule=schedule)
    if schedule == "dynamic":                 # a second batch uses a fresh counter
        assert population.evaluate_population(fake_evaluate, STRINGS[:5], schedule=schedule) == out[:5]
    queue.put((rank, out, calls))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("with_costs", [False, True])
def test_world_size_2_gloo(with_costs):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, with_costs, queue)) for r in range(2)]
    for p in procs:
        p.start()
    got = [queue.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = [fake_evaluate(s) for s in STRINGS]
    all_calls = []
    for rank, out, calls in got:
        assert out == want, rank              # every rank holds the full, identical result list
        all_calls += calls
    assert sorted(all_calls) == sorted(set(STRINGS))      # every unique individual exactly once
    sizes = sorted(len(c) for _, _, c in got)
    assert sizes[1] - sizes[0] <= 1                       # balanced


def test_work_queue_single_process():
    assert list(population.WorkQueue(5)) == [0, 1, 2, 3, 4]
    assert list(population.WorkQueue(0)) == []
    got = population.evaluate_population(fake_evaluate, STRINGS, schedule="dynamic")
    assert got == [fake_evaluate(s) for s in STRINGS]
    with pytest.raises(ValueError):
        population.evaluate_population(fake_evaluate, STRINGS, schedule="guided")


def test_concurrent_evaluators_share_the_jobs():
    import threading
    import time
    seen = {0: [], 1: [], 2: []}

    def make(k):
        def evaluate(s):
            seen[k].append((s, threading.get_ident()))
            time.sleep(0.005)                   # long enough for the other threads to take jobs too
            return fake_evaluate(s)
        return evaluate

    want = [fake_evaluate(s) for s in STRINGS]
    for schedule in ("static", "dynamic"):
        for k in seen:
            seen[k].clear()
        got = population.evaluate_population([make(0), make(1), make(2)], STRINGS, schedule=schedule)
        assert got == want
        calls = [s for k in seen for s, _ in seen[k]]
        assert sorted(calls) == sorted(set(STRINGS))                 # every unique individual exactly once
        assert len({t for k in seen for _, t in seen[k]}) >= 2       # on more than one thread

    def broken(s):
        raise RuntimeError("boom")
    with pytest.raises(RuntimeError):
        population.evaluate_population([broken, broken], STRINGS)


def test_world_size_2_gloo_dynamic_schedule():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, True, queue, "dynamic")) for r in range(2)]
    for p in procs:
        p.start()
    got = [queue.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = [fake_evaluate(s) for s in STRINGS]
    all_calls = []
    for rank, out, calls in got:
        assert out == want, rank
        all_calls += calls
    assert sorted(all_calls) == sorted(set(STRINGS))      # every unique individual exactly once
    by_rank = {rank: len(calls) for rank, _, calls in got}
    assert by_rank[1] > by_rank[0]                        # the fast rank took more of the queue


def test_world_size_2_gloo_dynamic_schedule_with_two_evaluators_per_rank():
    """The bench's configuration: ranks take individuals from the shared counter, two host threads each."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, True, queue, "dynamic", 2)) for r in range(2)]
    for p in procs:
        p.start()
    got = [queue.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = [fake_evaluate(s) for s in STRINGS]
    all_calls = []
    for rank, out, calls in got:
        assert out == want, rank
        all_calls += calls
    assert sorted(all_calls) == sorted(set(STRINGS))      # every unique individual exactly once


def test_batched_map_fitness_conventions():
    class Gen:
        def generate_and_evaluate(self, expression, storages, lo, hi, solver, infinity=1e100, evaluation_samples=1):
            return fake_evaluate(expression)

    class Gram:
        def compile(self, s):
            if s == "broken":
                raise SyntaxError(s)
            return s, None

    m = population.batched_map(Gen(), Gram(), 2, 4)
    out = m(None, ["ind_1", "broken", "ind_3"])
    t, rho, n = fake_evaluate("ind_1")
    assert out[0] == ((rho, t / n) if n < 1e100 else (rho, 1e100))
    assert out[1] == (1e100, 1e100)


def test_batched_map_objectives_and_size_guard():
    """optimization/program.py:386-453: single objective = (t,) or (sqrt(rho) sqrt(n),) when the solver did
    not converge; trees of more than 150 nodes get the infinity tuple without an evaluation."""
    import functools
    calls = []

    class Gen:
        def generate_and_evaluate(self, expression, storages, lo, hi, solver, infinity=1e100, evaluation_samples=1):
            calls.append(expression)
            return {"good": (12.0, 0.25, 6.0), "diverged": (3.0, 4.0, 1e100)}[expression]

    class Gram:
        def compile(self, s):
            return s, None

    class Optimizer:
        def evaluate_single_objective(self, individual, **kw):
            raise AssertionError("the mapped function is not called: the map evaluates the batch itself")

        def evaluate_multiple_objectives(self, individual, **kw):
            raise AssertionError

    big = "f(" * 151 + "x" + ")" * 151
    assert population.node_count(big) == 152 and population.node_count("f(1, (2, 1), g(x))") == 5
    m = population.batched_map(Gen(), Gram(), 2, 4)
    opt = Optimizer()
    single = m(functools.partial(opt.evaluate_single_objective, pset=None), ["good", "diverged", big])
    assert single == [(12.0,), (4.0 ** 0.5 * 1e100 ** 0.5,), (1e100,)]
    multi = m(opt.evaluate_multiple_objectives, ["good", "diverged", big])
    assert multi == [(0.25, 2.0), (4.0, 1e100), (1e100, 1e100)]
    assert big not in calls
    assert population.batched_map(Gen(), Gram(), 2, 4, objectives="single")(None, ["good"]) == [(12.0,)]

    class Tree(list):                     # a deap PrimitiveTree: len() = node count, str() = grammar string
        def __str__(self):
            return "good"
    assert m(None, [Tree(range(151)), Tree(range(150))]) == [(1e100, 1e100), (0.25, 2.0)]


def test_solo_gate_excludes_shared_holders():
    import threading
    import time
    gate = population.SoloGate()
    active, overlaps, log = [0], [0], []
    lock = threading.Lock()

    def reader():
        for _ in range(20):
            with gate.shared():
                with lock:
                    active[0] += 1
                time.sleep(0.001)
                with lock:
                    active[0] -= 1

    def writer():
        for _ in range(10):
            with gate.exclusive():
                with lock:
                    if active[0]:
                        overlaps[0] += 1
                    log.append("w")
                time.sleep(0.001)

    threads = [threading.Thread(target=reader) for _ in range(3)] + [threading.Thread(target=writer) for _ in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=30)
    assert overlaps[0] == 0 and len(log) == 20


def _failing_worker(rank, world, port, queue):
    sys.path.insert(0, REPO)
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def evaluate(s):
        if rank == 1 and s == "ind_3":
            raise RuntimeError("engine broke on rank 1")
        return fake_evaluate(s)

    try:
        population.evaluate_population(evaluate, STRINGS)
        queue.put((rank, "returned"))
    except RuntimeError as exc:
        queue.put((rank, str(exc)))
    dist.barrier()
    dist.destroy_process_group()


def test_a_failing_rank_does_not_leave_the_others_in_the_collective():
    """A rank whose evaluation raises still joins the result exchange; every rank raises afterwards."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_failing_worker, args=(r, 2, port, queue)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(queue.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[1] == "engine broke on rank 1"
    assert "failed on 1 other rank" in got[0]
