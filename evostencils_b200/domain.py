"""Domain decomposition of one 3D multigrid solve over the GPUs of a node (BASELINE.json configs[2]:
3D Poisson 512^3, V/W-cycles with red-black GS, 2/4/8 GPUs, halo exchange over NVLink).

The reference has no counterpart: its generated solvers run one block / one fragment with OpenMP
(/root/reference/example_problems/lib/domain_onePatch.knowledge:4-9,
lib/parallelization_pureOmp.knowledge:2-9; `mpi_enabled = false`).  Design:

* 1-D slabs along z (the slowest axis: a plane of n x pitch elements is contiguous, so a halo plane
  is sent as it lies in the field array -- no pack kernels).  Split points are multiples of
  2^(max_level - dd_min_level + 1) cells, so they coincide on every decomposed level (and one below) and restriction /
  prolongation need one ghost plane only.  Rank r owns the global planes (s_r, s_{r+1}] of a level.
* every op of the lowered cycle is executed by the rank-local slab workspace (mg_create_slab: the
  same CUDA kernels, geometry {nz, z0}); after each write of a solution-type field -- every colour
  pass of a red-black sweep, prolongation+correction -- and after a residual that a restriction
  reads, the two boundary planes are exchanged with the neighbours;
* levels below `dd_min_level` are agglomerated: the restricted right-hand side is all-gathered and
  every rank runs the coarse part of the cycle redundantly on a whole-grid workspace (no scatter);
* the residual norm is the square root of the rank-ordered sum of the local sums of squares.

Ranks are *virtual*: a process drives the ranks listed in `local_ranks`.  With torch.distributed
(one process per GPU, NCCL) that is its own rank and neighbours are reached by send/recv; in the
single-GPU parity tests one process drives all ranks and the exchange is a device copy, which
executes exactly the same schedule."""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

from . import lowering


# ----------------------------------------------------------------------------------------
# partition
# ----------------------------------------------------------------------------------------

def slab_splits(max_level: int, dd_min_level: int, world: int) -> List[int]:
    """Split points s_0 = 0 < s_1 < ... < s_world = 2^max_level (in fine cells / plane indices)."""
    units = 2 ** (dd_min_level - 1)          # split points must also exist one level below dd_min_level
    if units < world:
        raise ValueError(f"level {dd_min_level} has {2 * units} cells along z: too few for {world} slabs of >= 2")
    unit = 2 ** (max_level - dd_min_level + 1)
    return [unit * ((r * units) // world) for r in range(world)] + [2 ** max_level]


def owned_planes(level: int, max_level: int, splits: Sequence[int], rank: int) -> Tuple[int, int]:
    """First and last global interior plane rank `rank` owns on `level` (empty when last < first)."""
    shift = max_level - level
    if any(s % (1 << shift) for s in splits):
        raise ValueError("split points do not exist on this level")
    lo, hi = splits[rank] >> shift, splits[rank + 1] >> shift
    return lo + 1, min(hi, 2 ** level - 1)


def slab_ranges(min_level: int, max_level: int, splits: Sequence[int], rank: int) -> Dict[int, Tuple[int, int]]:
    return {level: owned_planes(level, max_level, splits, rank) for level in range(min_level, max_level + 1)}


# ----------------------------------------------------------------------------------------
# communication between (virtual) ranks: torch tensors, NCCL on the box / gloo in CPU tests
# ----------------------------------------------------------------------------------------

class SlabComm:
    """Neighbour exchange and gathers between the slabs of one decomposed solve.

    `local_ranks`: the virtual ranks this process drives.  `dist` = torch.distributed when the other
    ranks live in other processes (process rank == virtual rank), else None (all ranks local)."""

    def __init__(self, world: int, local_ranks: Sequence[int], dist=None, group=None):
        self.world = world
        self.local_ranks = list(local_ranks)
        self.dist = dist
        self.group = group
        if dist is None and sorted(self.local_ranks) != list(range(world)):
            raise ValueError("without torch.distributed one process must drive every rank")
        if dist is not None and len(self.local_ranks) != 1:
            raise ValueError("with torch.distributed a process drives exactly one rank")
        self.bytes_sent = 0

    @staticmethod
    def from_environment(group=None) -> "SlabComm":
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return SlabComm(dist.get_world_size(group), [dist.get_rank(group)], dist, group)
        return SlabComm(1, [0])

    def exchange_halos(self, views: Dict[int, object]):
        """views[rank] = tensor [nz][...] of that rank's slab; local planes 1 .. nz-2 are owned.
        Fills plane 0 from the lower neighbour's last owned plane and plane nz-1 from the upper
        neighbour's first owned plane.  Slabs without owned planes (nz == 2) take no part."""
        if self.dist is None:
            for r in range(self.world - 1):
                lo, hi = views[r], views[r + 1]
                if lo.shape[0] < 3 or hi.shape[0] < 3:
                    continue
                hi[0].copy_(lo[lo.shape[0] - 2])
                lo[lo.shape[0] - 1].copy_(hi[1])
                self.bytes_sent += 2 * lo[0].numel() * lo.element_size()
            return
        (r,) = self.local_ranks
        v = views[r]
        if v.shape[0] < 3:
            return
        ops = []
        P2POp, isend, irecv = self.dist.P2POp, self.dist.isend, self.dist.irecv
        if r > 0:
            ops.append(P2POp(isend, v[1], r - 1, self.group))
            ops.append(P2POp(irecv, v[0], r - 1, self.group))
        if r < self.world - 1:
            ops.append(P2POp(isend, v[v.shape[0] - 2], r + 1, self.group))
            ops.append(P2POp(irecv, v[v.shape[0] - 1], r + 1, self.group))
        self.bytes_sent += (len(ops) // 2) * v[0].numel() * v.element_size()
        if ops:
            for req in self.dist.batch_isend_irecv(ops):
                req.wait()

    def gather_planes(self, owned: Dict[int, object], counts: Sequence[int]):
        """owned[rank] = tensor [counts[rank]][...] (the rank's owned planes, possibly empty).
        -> the concatenation over ranks, on every rank (a tensor per local rank)."""
        import torch
        if self.dist is None:
            full = torch.cat([owned[r] for r in range(self.world)], dim=0)
            return {r: full for r in self.local_ranks}
        (r,) = self.local_ranks
        mine = owned[r].contiguous()
        m = max(counts)
        padded = torch.zeros((m,) + tuple(mine.shape[1:]), dtype=mine.dtype, device=mine.device)
        padded[: counts[r]] = mine
        out = [torch.empty_like(padded) for _ in range(self.world)]
        self.dist.all_gather(out, padded, group=self.group)
        self.bytes_sent += padded.numel() * padded.element_size()
        return {r: torch.cat([out[q][: counts[q]] for q in range(self.world)], dim=0)}

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier(group=self.group)

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier(group=self.group)

    def sum_in_rank_order(self, values: Dict[int, float]) -> float:
        """Deterministic: every rank adds the world's values in rank order."""
        import torch
        if self.dist is None:
            total = 0.0
            for r in range(self.world):
                total += values[r]
            return total
        (r,) = self.local_ranks
        device = "cuda" if self.dist.get_backend(self.group) == "nccl" else "cpu"
        mine = torch.tensor([values[r]], dtype=torch.float64, device=device)
        out = [torch.empty_like(mine) for _ in range(self.world)]
        self.dist.all_gather(out, mine, group=self.group)
        total = 0.0
        for t in out:
            total += float(t.item())
        return total


# ----------------------------------------------------------------------------------------
# the decomposed solve
# ----------------------------------------------------------------------------------------

def default_dd_min_level(min_level: int, world: int) -> int:
    """Coarsest decomposed level: every rank still owns >= 32 planes; the levels below are agglomerated
     -- an exchange costs tens of microseconds of latency whatever the plane size, and a
    replicated 65^3 level costs less than one exchange."""
    return max(min_level + 1, int(math.ceil(math.log2(32 * world))))


class _Local:
    def __init__(self, rank, slab, coarse):
        self.rank, self.slab, self.coarse = rank, slab, coarse


class DomainDecomposedSolver:
    """One 3D solve spread over `comm.world` slabs.  Same op vocabulary as the single-GPU engine
    (`lowering.Program.ops`, unfused), same kernels, same results bit for bit."""

    def __init__(self, problem, min_level: int, max_level: int, comm: SlabComm, dd_min_level: Optional[int] = None,
                 device_of=None, params: Optional[dict] = None):
        from .engine import EngineError, Workspace
        if problem.dimension != 3:
            raise EngineError("domain decomposition is implemented for 3D grids")
        self.problem, self.comm = problem, comm
        self.min_level, self.max_level = min_level, max_level
        self.dd_min_level = (min(max_level, default_dd_min_level(min_level, comm.world)) if dd_min_level is None
                             else dd_min_level)
        if not min_level < self.dd_min_level <= max_level:
            raise EngineError("dd_min_level must lie in (min_level, max_level]")
        self.base = self.dd_min_level - 1           # finest replicated level = extra coarse level of the slabs
        self.splits = slab_splits(max_level, self.dd_min_level, comm.world)
        self.params = dict(problem.parameters)
        self.params.update(params or {})
        device_of = device_of or (lambda rank: 0)
        self.locals: List[_Local] = []
        self._streams = {}
        self._previous_streams = {}
        for r in comm.local_ranks:
            ranges = slab_ranges(self.base, max_level, self.splits, r)
            slab = Workspace(3, self.base, max_level, problem.n_fields, problem.dtype, device_of(r), slab=ranges)
            coarse = Workspace(3, min_level, self.base, problem.n_fields, problem.dtype, device_of(r))
            for ws in (slab, coarse):
                for level in range(ws.min_level, ws.max_level + 1):
                    ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0, self.params)))
            coarse.set_coarse_solver(problem.solver.cgs, problem.solver.cgs_max_iterations,
                                     problem.solver.cgs_target_reduction)
            for ws in (slab, coarse):
                ws.exec_ops([])                     # one-time preparation (operator upload, kernel attributes)
                ws.set_stream(self._torch_stream(device_of(r)))
            self.locals.append(_Local(r, slab, coarse))
        self._pending_coarse: List[object] = []
        self.halo_exchanges = 0

    def _torch_stream(self, device: int) -> int:
        """Kernels, halo exchanges (NCCL on torch's current stream) and device copies share one stream,
        so a cycle is enqueued without any host synchronisation.  A dedicated non-default stream per
        device is made torch's current stream for the lifetime of the solver (the legacy default
        stream has handle 0 and cannot be handed to the engine)."""
        import torch
        if not torch.cuda.is_available():
            return 0
        if device not in self._streams:
            self._previous_streams[device] = torch.cuda.current_stream(device)
            self._streams[device] = torch.cuda.Stream(device)
            torch.cuda.set_stream(self._streams[device])
        return int(self._streams[device].cuda_stream)

    def close(self):
        self._sync()
        if self._previous_streams:
            import torch
            for device, stream in self._previous_streams.items():
                torch.cuda.set_stream(stream)
            self._previous_streams = {}
        for loc in self.locals:
            loc.slab.close()
            loc.coarse.close()
        self.locals = []

    # -- data ------------------------------------------------------------------------------------
    def plane_range(self, rank: int, level: int) -> Tuple[int, int]:
        """Global planes [z0, z0 + nz) held by `rank` on `level` (owned + ghost/boundary planes)."""
        zb, ze = owned_planes(level, self.max_level, self.splits, rank)
        return zb - 1, ze + 2

    def set_fields(self, rhs, u0):
        """rhs, u0: per field either a full host array [n][n][n] or a callable (z_lo, z_hi) -> slab array."""
        L = self.max_level
        for loc in self.locals:
            z_lo, z_hi = self.plane_range(loc.rank, L)
            for f in range(self.problem.n_fields):
                for kind, src in (("rhs", rhs[f]), ("sol", u0[f])):
                    loc.slab.upload(kind, L, f, src(z_lo, z_hi) if callable(src) else src[z_lo:z_hi])

    def gather_solution(self, field: int = 0):
        """Full finest-level solution on the host (tests; not part of the timed path)."""
        import numpy as np
        import torch
        L = self.max_level
        owned, counts = {}, []
        for r in range(self.comm.world):
            zb, ze = owned_planes(L, L, self.splits, r)
            counts.append(ze - zb + 1)
        for loc in self.locals:
            v = loc.slab.field_view("sol", L, field)
            owned[loc.rank] = v[1: v.shape[0] - 1]
        full = self.comm.gather_planes(owned, counts)[self.locals[0].rank]
        n = 2 ** L + 1
        out = np.zeros((n, n, n), dtype=self.locals[0].slab.np_dtype)
        out[1:-1] = full[:, :, :n].cpu().numpy()
        first = [loc for loc in self.locals if loc.rank == 0]
        last = [loc for loc in self.locals if loc.rank == self.comm.world - 1]
        if first:
            out[0] = first[0].slab.download("sol", L, field)[0]
        if last:
            out[-1] = last[0].slab.download("sol", L, field)[-1]
        return out

    # -- op interpreter ----------------------------------------------------------------------------
    def _sync(self):
        import torch
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def _flush_coarse(self):
        if self._pending_coarse:
            for loc in self.locals:
                loc.coarse.enqueue_ops(self._pending_coarse)
            self._pending_coarse = []

    def _slab_exec(self, op):
        for loc in self.locals:
            loc.slab.enqueue_ops([op])

    def _exchange(self, kind: str, level: int):
        for f in range(self.problem.n_fields):
            self.comm.exchange_halos({loc.rank: loc.slab.field_view(kind, level, f) for loc in self.locals})
        self.halo_exchanges += 1

    def run_ops(self, ops: Sequence[object]):
        import copy
        base, dd = self.base, self.dd_min_level
        for op in ops:
            kind = op.kind
            if kind in ("residual_restrict",):
                raise ValueError("pass the unfused op list: the decomposed restriction reads exchanged residual planes")
            if kind == "smooth":
                if op.level <= base:
                    self._pending_coarse.append(op)
                    continue
                self._flush_coarse()
                if op.colored:
                    for colour_pass in (1, 2):
                        half = copy.copy(op)
                        half.colour_pass = colour_pass
                        self._slab_exec(half)
                        self._exchange("sol", op.level)
                else:
                    self._slab_exec(op)
                    self._exchange("sol", op.level)
            elif kind == "residual":
                if op.level <= base:
                    self._pending_coarse.append(op)
                    continue
                self._flush_coarse()
                self._slab_exec(op)
                self._exchange("res", op.level)
            elif kind == "restrict":
                if op.level < base:
                    self._pending_coarse.append(op)
                    continue
                self._flush_coarse()
                self._slab_exec(op)
                if op.level == base:
                    self._rhs_to_coarse()
            elif kind == "zero":
                if op.level <= base:
                    self._pending_coarse.append(op)
                else:
                    self._flush_coarse()
                    self._slab_exec(op)
            elif kind == "prolong_add":
                if op.level <= base:
                    self._pending_coarse.append(op)
                    continue
                self._flush_coarse()
                if op.level == dd:
                    self._sol_from_coarse()
                self._slab_exec(op)
                self._exchange("sol", op.level)
            elif kind == "cgs":
                if op.level != self.min_level or op.inner:
                    raise ValueError("decomposed solves place the coarse-grid solver on the engine's min level")
                self._pending_coarse.append(op)
            else:
                raise ValueError(f"op kind '{kind}' is not available in a decomposed solve")
        self._flush_coarse()

    def _rhs_to_coarse(self):
        """All-gather the restricted right-hand side of level `base` into the replicated workspaces."""
        level = self.base
        counts = []
        for r in range(self.comm.world):
            zb, ze = owned_planes(level, self.max_level, self.splits, r)
            counts.append(max(0, ze - zb + 1))
        for f in range(self.problem.n_fields):
            owned = {}
            for loc in self.locals:
                v = loc.slab.field_view("rhs", level, f)
                owned[loc.rank] = v[1: 1 + counts[loc.rank]]
            full = self.comm.gather_planes(owned, counts)
            for loc in self.locals:
                loc.coarse.field_view("rhs", level, f)[1:-1].copy_(full[loc.rank])

    def _sol_from_coarse(self):
        """Slab copy of the replicated coarse solution of level `base` (local: every rank has all of it)."""
        level = self.base
        for loc in self.locals:
            z_lo, z_hi = self.plane_range(loc.rank, level)
            for f in range(self.problem.n_fields):
                loc.slab.field_view("sol", level, f).copy_(loc.coarse.field_view("sol", level, f)[z_lo:z_hi])

    # -- outer iteration ---------------------------------------------------------------------------
    def residual_norm(self) -> float:
        partial = {}
        for loc in self.locals:
            v = loc.slab.residual_norm(self.max_level)
            partial[loc.rank] = v * v
        return math.sqrt(self.comm.sum_in_rank_order(partial))

    def time_cycles(self, ops: Sequence[object], cycles: int, warmup: int = 1):
        """`cycles` cycles back to back between device-synchronised barriers -> (ms per cycle on this
        rank, record for the bench line).  The caller takes the max over ranks."""
        import time
        for _ in range(warmup):
            self.run_ops(ops)
        self.comm.bytes_sent = 0
        self.halo_exchanges = 0
        self.comm.barrier()
        self._sync()
        t0 = time.perf_counter()
        for _ in range(cycles):
            self.run_ops(ops)
        self._sync()
        self.comm.barrier()
        ms = (time.perf_counter() - t0) * 1e3 / cycles
        return ms, {"driver": "op-by-op host driver, NCCL send/recv halo exchange (no graph)",
                    "halo_bytes_per_cycle_rank0": self.comm.bytes_sent / cycles,
                    "halo_exchanges_per_cycle": self.halo_exchanges / cycles,
                    "dd_min_level": self.dd_min_level, "splits": self.splits}

    def time_cycles(self, ops: Sequence[object], cycles: int, warmup: int = 1):
        """`cycles` cycles back to back between device-synchronised barriers -> (ms per cycle, record)."""
        import time
        for _ in range(warmup):
            self.run_ops(ops)
        self.comm.bytes_sent = 0
        self.halo_exchanges = 0
        self.comm.barrier()
        self._sync()
        t0 = time.perf_counter()
        for _ in range(cycles):
            self.run_ops(ops)
        self._sync()
        self.comm.barrier()
        ms = (time.perf_counter() - t0) * 1e3 / cycles
        return ms, {"driver": "op-by-op host driver, NCCL send/recv halo exchange (no graph)",
                    "halo_bytes_per_cycle_rank0": self.comm.bytes_sent / cycles,
                    "halo_exchanges_per_cycle": self.halo_exchanges / cycles,
                    "dd_min_level": self.dd_min_level, "splits": self.splits}

    def solve(self, ops: Sequence[object], tol: float, max_its: int):
        """Outer iteration of the generated solver (same conventions as mg_solve)
        -> (history, converged, milliseconds)."""
        import time
        self._sync()
        t0 = time.perf_counter()
        history = [self.residual_norm()]
        converged = False
        for _ in range(max_its):
            self.run_ops(ops)
            r = self.residual_norm()
            history.append(r)
            if not math.isfinite(r):
                break
            if r < tol * history[0]:
                converged = True
                break
        self._sync()
        return history, converged, (time.perf_counter() - t0) * 1e3


def make_solver(problem, min_level: int, max_level: int, comm: SlabComm, device: int = 0, dd_min_level=None):
    """The decomposed solver for this process' rank(s) on `device`."""
    return DomainDecomposedSolver(problem, min_level, max_level, comm, dd_min_level=dd_min_level,
                                  device_of=lambda r: device)


# ----------------------------------------------------------------------------------------
# peer-to-peer transport: halo planes stored straight into the neighbour's memory (csrc/mg_dd.cuh)
# ----------------------------------------------------------------------------------------

def encode_for_slabs(slab, coarse, ops):
    """MgOp array of the unfused cycle for mg_dd_build_cycle: smoother tables of ops on levels up to the
    coarsest slab level are registered with (and numbered by) the coarse workspace, the others with the slab."""
    from .engine import MgOp
    flat = []
    base = slab.min_level
    for op in ops:
        ws = coarse if op.level <= base else slab
        flat.extend(ws.encode([op]))
    return (MgOp * len(flat))(*flat)


class PeerSlabSolver:
    """One 3D solve over `world` z-slabs whose ranks exchange halo planes, the agglomerated right-hand side
    and the partial norms by peer-to-peer stores / loads inside the engine's own kernels (NVLink between the
    GPUs of a node; CUDA IPC maps the neighbours' arrays) -- torch.distributed only ships the memory
    descriptors once at set-up.  A cycle of a rank is one CUDA-graph launch; the host reads one double per
    cycle.  Ranks are virtual like in `DomainDecomposedSolver`: with `dist` a process drives its own rank, in
    the single-GPU tests one process drives all of them (same kernels, same handshakes, plain pointers).
    Results are bitwise those of the single-domain engine."""

    def __init__(self, problem, min_level: int, max_level: int, world: int, local_ranks: Sequence[int], dist=None,
                 group=None, dd_min_level: Optional[int] = None, device_of=None, params: Optional[dict] = None):
        from .engine import EngineError, Workspace
        if problem.dimension != 3 or problem.n_fields != 1 or problem.dtype != "f64":
            raise EngineError("the peer-to-peer decomposition is implemented for scalar fp64 3D problems")
        if dist is not None and len(local_ranks) != 1:
            raise ValueError("with torch.distributed a process drives exactly one rank")
        if dist is None and sorted(local_ranks) != list(range(world)):
            raise ValueError("without torch.distributed one process must drive every rank")
        self.problem, self.world, self.dist, self.group = problem, world, dist, group
        self.min_level, self.max_level = min_level, max_level
        self.dd_min_level = (min(max_level, default_dd_min_level(min_level, world)) if dd_min_level is None
                             else dd_min_level)
        if not min_level < self.dd_min_level <= max_level:
            raise EngineError("dd_min_level must lie in (min_level, max_level]")
        self.base = self.dd_min_level - 1
        self.splits = slab_splits(max_level, self.dd_min_level, world)
        self.params = dict(problem.parameters)
        self.params.update(params or {})
        device_of = device_of or (lambda rank: 0)
        self.locals: List[_Local] = []
        for r in local_ranks:
            ranges = slab_ranges(self.base, max_level, self.splits, r)
            slab = Workspace(3, self.base, max_level, 1, "f64", device_of(r), slab=ranges)
            coarse = Workspace(3, min_level, self.base, 1, "f64", device_of(r))
            for ws in (slab, coarse):
                for level in range(ws.min_level, ws.max_level + 1):
                    ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0, self.params)))
            coarse.set_coarse_solver(problem.solver.cgs, problem.solver.cgs_max_iterations,
                                     problem.solver.cgs_target_reduction)
            slab.dd_init(r, world, coarse)
            self.locals.append(_Local(r, slab, coarse))
        self._connect()
        self._cycles: Dict[int, int] = {}

    # -- set-up: every rank learns where its peers' arrays live ----------------------------------------
    def _connect(self):
        ipc = self.dist is not None
        mine = {}
        for loc in self.locals:
            table = {-1: loc.slab.dd_describe(-1, ipc)}
            for level in range(self.base, self.max_level + 1):
                table[level] = loc.slab.dd_describe(level, ipc)
            mine[loc.rank] = table
        if ipc:
            gathered = [None] * self.world
            self.dist.all_gather_object(gathered, mine, group=self.group)
            everyone = {}
            for part in gathered:
                everyone.update(part)
        else:
            everyone = mine
        for loc in self.locals:
            for q in range(self.world):
                loc.slab.dd_attach(q, -1, *everyone[q][-1])
                loc.slab.dd_attach(q, self.base, *everyone[q][self.base])
            for q in (loc.rank - 1, loc.rank + 1):
                if 0 <= q < self.world:
                    for level in range(self.base + 1, self.max_level + 1):
                        loc.slab.dd_attach(q, level, *everyone[q][level])
        self.barrier()

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier(group=self.group)

    def close(self):
        for loc in self.locals:
            loc.slab.synchronize()
        self.barrier()                      # nobody unmaps memory a peer's kernel may still touch
        for loc in self.locals:
            loc.coarse.close()              # first: it runs on the slab workspace's stream (mg_dd_init)
            loc.slab.close()
        self.locals = []

    # -- data --------------------------------------------------------------------------------------------
    def plane_range(self, rank: int, level: int) -> Tuple[int, int]:
        zb, ze = owned_planes(level, self.max_level, self.splits, rank)
        return zb - 1, ze + 2

    def set_fields(self, rhs, u0):
        """rhs, u0: per field either a full host array [n][n][n] or a callable (z_lo, z_hi) -> slab array.
        The second ghost planes of the solution (read by the one-pass smoother) are filled by an exchange."""
        L = self.max_level
        for loc in self.locals:
            z_lo, z_hi = self.plane_range(loc.rank, L)
            for kind, src in (("rhs", rhs[0]), ("sol", u0[0])):
                loc.slab.upload(kind, L, 0, src(z_lo, z_hi) if callable(src) else src[z_lo:z_hi])
        self.barrier()
        for loc in self.locals:
            loc.slab.dd_exchange("sol", L, 2)
        for loc in self.locals:
            loc.slab.synchronize()

    def local_solution(self, rank: int):
        """Host copy [nz][n][n] of a local rank's finest-level slab (ghost planes included)."""
        (loc,) = [l for l in self.locals if l.rank == rank]
        return loc.slab.download("sol", self.max_level, 0)

    def gather_solution(self):
        """Full finest-level solution on the host; needs every rank local (tests)."""
        import numpy as np
        L = self.max_level
        n = 2 ** L + 1
        out = np.zeros((n, n, n))
        for loc in self.locals:
            z_lo, z_hi = self.plane_range(loc.rank, L)
            slab = loc.slab.download("sol", L, 0)
            zb, ze = owned_planes(L, L, self.splits, loc.rank)
            out[zb:ze + 1] = slab[zb - z_lo: ze + 1 - z_lo]
            if loc.rank == 0:
                out[0] = slab[0]
            if loc.rank == self.world - 1:
                out[-1] = slab[-1]
        return out

    # -- execution ---------------------------------------------------------------------------------------
    def cycle_handle(self, ops) -> int:
        key = id(ops)
        if key not in self._cycles:
            handle = None
            for loc in self.locals:
                h = loc.slab.dd_build_cycle(encode_for_slabs(loc.slab, loc.coarse, ops))
                assert handle is None or handle == h
                handle = h
                loc.slab.dd_prepare(h)
            self._cycles[key] = handle
            self._keep = getattr(self, "_keep", []) + [ops]       # ids stay unique while the list lives
            self.barrier()                   # every rank's graphs exist before anybody's kernels wait for a peer
        return self._cycles[key]

    def launch(self, ops, n: int = 1):
        handle = self.cycle_handle(ops)
        for loc in self.locals:
            loc.slab.dd_launch(handle, n)

    def read_norm(self) -> float:
        values = [loc.slab.dd_read_norm() for loc in self.locals]
        assert all(v == values[0] or (v != v and values[0] != values[0]) for v in values), values
        return values[0]

    def residual_norm(self) -> float:
        for loc in self.locals:
            loc.slab.dd_norm()
        return self.read_norm()

    def solve(self, ops: Sequence[object], tol: float, max_its: int):
        """Outer iteration of the generated solver (same conventions as mg_solve) -> (history, converged, ms)."""
        import time
        handle = self.cycle_handle(ops)
        t0 = time.perf_counter()
        history = [self.residual_norm()]
        converged = False
        for _ in range(max_its):
            for loc in self.locals:
                loc.slab.dd_launch(handle, 1)
            r = self.read_norm()
            history.append(r)
            if not math.isfinite(r):
                break
            if r < tol * history[0]:
                converged = True
                break
        return history, converged, (time.perf_counter() - t0) * 1e3

    def time_cycles(self, ops: Sequence[object], cycles: int, warmup: int = 1):
        """`cycles` graph launches back to back -> (device ms per cycle on this process, record)."""
        handle = self.cycle_handle(ops)
        launches0 = sum(loc.slab.launch_count() for loc in self.locals)
        for loc in self.locals:
            loc.slab.dd_launch(handle, warmup)
        self.read_norm()
        self.barrier()
        for loc in self.locals:
            loc.slab.mark(0)
            loc.slab.dd_launch(handle, cycles)
            loc.slab.mark(1)
        self.read_norm()
        ms = max(loc.slab.elapsed() for loc in self.locals) / cycles
        self.barrier()
        kernels = (sum(loc.slab.launch_count() for loc in self.locals) - launches0) / (cycles + warmup) / len(self.locals)
        return ms, {"driver": "one CUDA graph per cycle and rank; halo planes, gathered right-hand side and partial "
                              "norms move by peer-to-peer stores / loads inside the engine's kernels (no NCCL call)",
                    "kernels_per_cycle_and_rank": kernels, "dd_min_level": self.dd_min_level, "splits": self.splits,
                    "timing": "CUDA events on the rank's stream around the launches"}


def make_solver(problem, min_level: int, max_level: int, comm: SlabComm, device: int = 0, dd_min_level=None,
                transport: str = "peer"):
    """The decomposed solver for this process' rank(s) on `device`: `peer` = PeerSlabSolver (the product
    path), `nccl` = the op-by-op driver with torch.distributed send/recv (kept for comparison)."""
    if transport == "peer":
        return PeerSlabSolver(problem, min_level, max_level, comm.world, comm.local_ranks, comm.dist, comm.group,
                              dd_min_level=dd_min_level, device_of=lambda r: device)
    return DomainDecomposedSolver(problem, min_level, max_level, comm, dd_min_level=dd_min_level,
                                  device_of=lambda r: device)
