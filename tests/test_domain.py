"""Domain decomposition: partition arithmetic and the halo-exchange /
gather plumbing on CPU (gloo, world_size 2); the decomposed solve itself against the single-domain
engine on one GPU with virtual ranks (bitwise)."""
import os
import socket
import sys

import numpy as np
import pytest

from evostencils_b200 import domain

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_splits_cover_all_interior_planes_on_every_level():
    for world, dd, top in ((2, 3, 6), (4, 4, 7), (8, 5, 9), (3, 3, 5)):
        splits = domain.slab_splits(top, dd, world)
        assert splits[0] == 0 and splits[-1] == 2 ** top and sorted(splits) == splits
        for level in range(dd - 1, top + 1):
            planes = []
            for r in range(world):
                zb, ze = domain.owned_planes(level, top, splits, r)
                planes += list(range(zb, ze + 1))
                if level >= dd:
                    assert ze - zb + 1 >= 1
            assert planes == list(range(1, 2 ** level))          # every interior plane exactly once, in order
    with pytest.raises(ValueError):
        domain.slab_splits(6, 2, 4)


def test_nested_splits_make_transfers_local():
    """Coarse planes a rank needs for prolongation / fine residual planes it needs for restriction
    lie inside its owned + ghost range (ghost width 1)."""
    world, dd, top = 4, 4, 7
    splits = domain.slab_splits(top, dd, world)
    for level in range(dd, top + 1):
        for r in range(world):
            zb, ze = domain.owned_planes(level, top, splits, r)
            cb, ce = domain.owned_planes(level - 1, top, splits, r)
            for z in range(zb, ze + 1):                            # prolongation of fine plane z
                assert cb - 1 <= z // 2 and (z + 1) // 2 <= ce + 1
            for Z in range(cb, ce + 1):                            # restriction to coarse plane Z
                assert zb - 1 <= 2 * Z - 1 and 2 * Z + 1 <= ze + 1


def _slab(rank, nz):
    import torch
    t = torch.zeros((nz, 3, 4), dtype=torch.float64)
    for k in range(1, nz - 1):
        t[k] = 100.0 * rank + k
    return t


def test_loopback_exchange_and_gather():
    import torch
    comm = domain.SlabComm(3, [0, 1, 2])
    views = {0: _slab(0, 5), 1: _slab(1, 4), 2: _slab(2, 6)}
    comm.exchange_halos(views)
    assert torch.all(views[0][4] == 101.0) and torch.all(views[1][0] == 3.0)
    assert torch.all(views[1][3] == 201.0) and torch.all(views[2][0] == 102.0)
    assert torch.all(views[0][0] == 0.0) and torch.all(views[2][5] == 0.0)     # global boundary planes untouched
    full = comm.gather_planes({r: views[r][1:-1] for r in range(3)}, [3, 2, 4])
    assert full[0].shape[0] == 9 and float(full[0][3, 0, 0]) == 101.0
    assert comm.sum_in_rank_order({0: 1.0, 1: 2.0, 2: 4.0}) == 7.0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, queue):
    sys.path.insert(0, REPO)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from evostencils_b200 import domain as d
    comm = d.SlabComm.from_environment()
    nz = 5 if rank == 0 else 4
    view = _slab(rank, nz)
    comm.exchange_halos({rank: view})
    counts = [3, 2]
    full = comm.gather_planes({rank: view[1:-1]}, counts)[rank]
    total = comm.sum_in_rank_order({rank: float(rank + 1) * 0.1})
    queue.put((rank, view.numpy().copy(), full.numpy().copy(), total))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo_exchange():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, queue)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in procs:
        rank, view, full, total = queue.get(timeout=120)
        got[rank] = (view, full, total)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.all(got[0][0][4] == 101.0) and np.all(got[1][0][0] == 3.0)
    assert np.all(got[0][0][0] == 0.0) and np.all(got[1][0][3] == 0.0)
    for r in (0, 1):
        assert got[r][1].shape[0] == 5
        assert [float(got[r][1][k, 0, 0]) for k in range(5)] == [1.0, 2.0, 3.0, 101.0, 102.0]
        assert got[r][2] == 0.1 + 0.2


@pytest.mark.gpu
@pytest.mark.parametrize("world,lo,hi", [(2, 2, 5), (4, 2, 6), (3, 2, 5)])
def test_decomposed_solve_is_bitwise_the_single_domain_solve(world, lo, hi):
    """Virtual ranks on one GPU, device-copy halo exchange: same schedule as the NCCL path."""
    from evostencils_b200 import lowering, problems
    from evostencils_b200.engine import Workspace
    from evostencils_b200.grammar import Grammar, v_cycle_string
    problem = problems.POISSON_3D
    grammar = Grammar(problem, lo, hi)
    expression, _ = grammar.compile(v_cycle_string(hi - lo, 2, 1, 23, "red_black"))
    program = lowering.lower(expression, lo, hi)
    us, fs = problem.initial_fields(hi, {})
    rng = np.random.default_rng(5)
    fs = [f + 0 for f in fs]
    fs[0][1:-1, 1:-1, 1:-1] += rng.standard_normal(fs[0][1:-1, 1:-1, 1:-1].shape)

    ws = Workspace(3, lo, hi, 1, "f64", 0)
    for level in range(lo, hi + 1):
        ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    ws.set_coarse_solver(problem.solver.cgs, problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
    ws.upload("rhs", hi, 0, fs[0])
    ws.upload("sol", hi, 0, us[0])
    single_hist = [ws.residual_norm()]
    for _ in range(3):
        ws.exec_ops(program.ops, fused=False)
        single_hist.append(ws.residual_norm())
    want = ws.download("sol", hi, 0)
    ws.close()

    comm = domain.SlabComm(world, list(range(world)))
    import math
    deepest = max(lo + 1, int(math.ceil(math.log2(world))) + 1)         # decompose as many levels as possible
    dd = domain.DomainDecomposedSolver(problem, lo, hi, comm, dd_min_level=deepest)
    dd.set_fields(fs, us)
    hist, converged, ms = dd.solve(program.ops, 0.0, 3)
    got = dd.gather_solution()
    assert dd.halo_exchanges > 0
    dd.close()
    assert np.array_equal(got, want)
    assert np.allclose(hist, single_hist, rtol=1e-12, atol=0)


@pytest.mark.gpu
def test_decomposed_solve_with_the_specialised_3d_kernels(monkeypatch):
    """Levels 6 and 7 run the 2.5-D blocked 7-point kernels on slabs (tail kernel restricted to
    levels <= 3 so that the grid kernels see slab geometry on the levels above)."""
    monkeypatch.setenv("MG_TAIL_LEVEL", "3")
    test_decomposed_solve_is_bitwise_the_single_domain_solve(2, 2, 7)
    test_decomposed_solve_is_bitwise_the_single_domain_solve(4, 3, 7)


# ---- peer-to-peer transport (csrc/mg_dd.cuh): one CUDA graph per cycle and rank ---------------------------------

def _single_domain(problem, program, lo, hi, fs, us, cycles):
    from evostencils_b200 import lowering
    from evostencils_b200.engine import Workspace
    ws = Workspace(3, lo, hi, 1, "f64", 0)
    for level in range(lo, hi + 1):
        ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    ws.set_coarse_solver(problem.solver.cgs, problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
    ws.upload("rhs", hi, 0, fs[0])
    ws.upload("sol", hi, 0, us[0])
    hist = [ws.residual_norm()]
    for _ in range(cycles):
        ws.exec_ops(program.ops, fused=False)
        hist.append(ws.residual_norm())
    u = ws.download("sol", hi, 0)
    ws.close()
    return hist, u


def _case(lo, hi, seed=5):
    from evostencils_b200 import lowering, problems
    from evostencils_b200.grammar import Grammar, v_cycle_string
    problem = problems.POISSON_3D
    expression, _ = Grammar(problem, lo, hi).compile(v_cycle_string(hi - lo, 2, 1, 23, "red_black"))
    program = lowering.lower(expression, lo, hi)
    us, fs = problem.initial_fields(hi, {})
    fs = [f + 0 for f in fs]
    fs[0][1:-1, 1:-1, 1:-1] += np.random.default_rng(seed).standard_normal(fs[0][1:-1, 1:-1, 1:-1].shape)
    return problem, program, us, fs


@pytest.mark.gpu
@pytest.mark.parametrize("world,lo,hi,dd", [(2, 2, 5, 3), (4, 2, 6, 4), (3, 2, 5, 3), (2, 2, 6, 6)])
def test_peer_transport_is_bitwise_the_single_domain_solve(world, lo, hi, dd):
    """Virtual ranks on one GPU: the ranks' kernels run on different streams of the same device and hand
    halo planes, the gathered right-hand side and the partial norms to each other through the flag
    handshakes of mg_dd.cuh (plain pointers instead of CUDA IPC mappings); decomposition down to level `dd`.
    (Graph replays only: one host thread drives all ranks here, and direct launches synchronise a rank's
    stream now and then -- the tail program ring -- before the peers' work is enqueued.  Direct launches
    are covered by the two-process test.)"""
    problem, program, us, fs = _case(lo, hi)
    want_hist, want = _single_domain(problem, program, lo, hi, fs, us, 3)
    solver = domain.PeerSlabSolver(problem, lo, hi, world, list(range(world)), dd_min_level=dd)
    solver.set_fields(fs, us)
    hist, converged, ms = solver.solve(program.ops, 0.0, 3)
    got = solver.gather_solution()
    solver.close()
    assert np.array_equal(got, want)
    assert np.allclose(hist, want_hist, rtol=1e-12, atol=0)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3])
def test_peer_transport_with_the_one_pass_kernels_on_slabs(world, monkeypatch):
    """129^3 with the thresholds of k_rbgs3d / k_resrestrict3d lowered to 65: levels 6 and 7 are slabs with two
    ghost planes, one exchange per sweep, fused residual + restriction on the slabs -- still bitwise the
    single-domain result (which runs the same kernels on the whole grid)."""
    monkeypatch.setenv("MG_RBGS3D_MIN_N", "65")
    monkeypatch.setenv("MG_RESTRICT3D_MIN_N", "65")
    lo, hi = 2, 7
    problem, program, us, fs = _case(lo, hi, seed=6)
    want_hist, want = _single_domain(problem, program, lo, hi, fs, us, 2)
    solver = domain.PeerSlabSolver(problem, lo, hi, world, list(range(world)), dd_min_level=6)
    solver.set_fields(fs, us)
    hist, converged, ms = solver.solve(program.ops, 0.0, 2)
    got = solver.gather_solution()
    ms_cycle, info = solver.time_cycles(program.ops, 2, warmup=1)
    solver.close()
    assert np.array_equal(got, want)
    assert np.allclose(hist, want_hist, rtol=1e-12, atol=0)
    # V(2,1): 3 sweeps + fused restriction + prolongation per slab level, far fewer launches than exchanges + ops
    assert info["kernels_per_cycle_and_rank"] < 60 and ms_cycle > 0


def _peer_worker(rank, world, port, queue, graph):
    sys.path.insert(0, REPO)
    os.environ["MG_DD_GRAPH"] = graph
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from evostencils_b200 import domain as d
    lo, hi = 2, 7
    problem, program, us, fs = _case(lo, hi, seed=7)
    os.environ["MG_RBGS3D_MIN_N"] = "65"
    os.environ["MG_RESTRICT3D_MIN_N"] = "65"
    solver = d.PeerSlabSolver(problem, lo, hi, world, [rank], dist, dd_min_level=6, device_of=lambda r: rank)
    solver.set_fields(fs, us)
    hist, converged, ms = solver.solve(program.ops, 0.0, 3)
    z_lo, z_hi = solver.plane_range(rank, hi)
    mine = solver.local_solution(rank)
    solver.close()
    queue.put((rank, hist, z_lo, z_hi, mine))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("graph", ["1", "0"])
def test_peer_transport_between_two_processes_over_cuda_ipc(graph):
    """Two processes, two GPUs (run with `gpurun --gpus 2`; skipped on a single-GPU box): the neighbours'
    arrays are mapped with cudaIpcOpenMemHandle and the halo planes travel over NVLink."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    lo, hi = 2, 7
    os.environ["MG_RBGS3D_MIN_N"] = "65"
    os.environ["MG_RESTRICT3D_MIN_N"] = "65"
    try:
        problem, program, us, fs = _case(lo, hi, seed=7)
        want_hist, want = _single_domain(problem, program, lo, hi, fs, us, 3)
    finally:
        del os.environ["MG_RBGS3D_MIN_N"], os.environ["MG_RESTRICT3D_MIN_N"]
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_peer_worker, args=(r, 2, port, queue, graph)) for r in range(2)]
    for p in procs:
        p.start()
    got = [queue.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, hist, z_lo, z_hi, mine in got:
        assert np.allclose(hist, want_hist, rtol=1e-12, atol=0), rank
        assert np.array_equal(mine[1:-1], want[z_lo + 1:z_hi - 1]), rank
    assert got[0][1] == got[1][1]                    # the rank-ordered norm is bitwise the same on both ranks
