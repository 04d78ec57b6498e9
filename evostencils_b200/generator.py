"""B200ProgramGenerator -- drop-in for the reference's `ProgramGenerator` on the
fitness-evaluation path.

Mirrors the members `evostencils.optimization.program.Optimizer` touches
(/root/reference/evostencils/code_generation/exastencils.py):
  constructor side products  :112-194      generate_storage            :586-592
  initialize_code_generation :445-466      generate_and_evaluate       :485-537
  generate_cycle_function    :318-336      reinitialize                :196-215
  parse_output conventions   :539-584      evaluate (sample averaging) :417-443
Instead of emitting ExaSlang, running the ExaStencils compiler twice, `make`, and the
generated binary, an individual is lowered to an op list and executed by the CUDA
engine through the C ABI (engine.py / include/mgengine.h).  Same call, same result
triple `(time_to_solution_ms, convergence_factor, number_of_iterations)`, same
sentinel behaviour: a bad individual never raises, it returns `(infinity,)*3`.
"""
from __future__ import annotations

import contextlib
import math
import os
import time
from typing import Dict, List, Optional

import numpy as np

from . import ir, lowering, problems, stencils
from .engine import EngineError, Workspace


class CycleStorage:
    """Field names of one level (exastencils.py:14-26); opaque to the optimizer."""

    def __init__(self, problem: problems.ProblemSpec, grids, max_level):
        self.grid = grids
        level = grids[0].level
        self.solution = [f"{f}@{level}" if level == max_level else f"gen_error_{f}@{level}" for f in problem.field_names]
        self.rhs = [f"{r}@{level}" for r in problem.rhs_names]
        self.residual = [f"gen_residual_{f}@{level}" for f in problem.field_names]
        self.correction = [f"gen_error_{f}@{level}" for f in problem.field_names]


def _parse_knowledge(path):
    dimension = min_level = max_level = None
    with open(path) as fh:
        for line in fh:
            tokens = line.split("=")
            lhs = tokens[0].strip(" \n\t")
            if lhs == "dimensionality":
                dimension = int(tokens[1].strip(" \n\t"))
            elif lhs == "minLevel":
                min_level = int(tokens[1].strip(" \n\t"))
            elif lhs == "maxLevel":
                max_level = int(tokens[1].strip(" \n\t"))
    return dimension, min_level, max_level


def _parse_config_name(path):
    with open(path) as fh:
        for line in fh:
            tokens = line.split("=")
            if tokens[0].strip(" \n\t") == "configName":
                return tokens[1].strip(' \n\t"')
    return None


def summarise_history(history, n, converged, infinity=1e100, solver_iteration_limit=None):
    """History of residual norms -> (convergence_factor, n_iterations) with the conventions of
    `parse_output` (exastencils.py:539-584): geometric mean of the per-iteration factors,
    NaN/Inf factors count as sqrt(infinity); hitting the iteration limit -> n = infinity."""
    rho_inf = math.sqrt(infinity)
    factors = []
    count = 0
    for a, b in zip(history[:-1], history[1:]):
        rho = b / a if a != 0 else float("nan")
        if math.isinf(rho) or math.isnan(rho):
            factors.append(rho_inf)
        else:
            factors.append(rho)
            count += 1
    if count > 0:
        factor = 1.0
        exponent = 1.0 / len(factors)
        for rho in factors:
            factor *= math.pow(rho, exponent)
    else:
        factor = infinity
    iterations = float(n)
    if not converged or n == 0:
        iterations = infinity
    if solver_iteration_limit is not None and iterations >= solver_iteration_limit:
        iterations = infinity
    return factor, iterations


def summarise_bicgstab(history, converged, infinity=1e100, solver_iteration_limit=None):
    """Output contract of the hand-written outer solver (Helmholtz/...exa3:192-199): one
    "convergence factor" line = |curRes| / |initRes|; the printed iteration number is the 0-based
    loop counter `curStep`, and `parse_output` maps 0 and the limit message to infinity."""
    factor = history[-1] / history[0] if history[0] != 0 else float("nan")
    if math.isinf(factor) or math.isnan(factor):
        factor = math.sqrt(infinity)
    steps = len(history) - 2                       # curStep of the iteration that met the tolerance
    iterations = float(steps) if converged and steps > 0 else infinity
    if len(history) == 1:
        return infinity, infinity
    if solver_iteration_limit is not None and iterations >= solver_iteration_limit:
        iterations = infinity
    return factor, iterations


class B200ProgramGenerator:
    def __init__(self, absolute_compiler_path: str = "", base_path: str = "", settings_path: str = "",
                 knowledge_path: str = "", platform_path: str = "", mpi_rank=0, solution_equations=None,
                 cycle_name="gen_mgCycle", model_based_estimation=False, evaluation_timeout=300,
                 code_generation_timeout=300, c_compiler_timeout=120, solver_iteration_limit=None,
                 problem: Optional[problems.ProblemSpec] = None, min_level=None, max_level=None,
                 device: Optional[int] = None, engine_min_level: Optional[int] = None,
                 resident_inputs: bool = False, prefetch_inputs: bool = False):
        """The first twelve arguments are the reference's (exastencils.py:40-43); the compiler
        and platform paths are accepted and ignored.  `problem` (or the `configName` of the
        settings file) selects the problem definition; levels come from the knowledge file
        when it exists, else from the problem defaults."""
        self._average_generation_time = 0
        self._counter = 0
        self.timeout_evaluate = evaluation_timeout
        self._solver_iteration_limit = solver_iteration_limit
        self._mpi_rank = mpi_rank
        self._cycle_name = cycle_name
        self._base_path = base_path
        self._platform_path = platform_path
        self._absolute_compiler_path = absolute_compiler_path
        name = None
        k_dim = k_min = k_max = None
        if settings_path and os.path.exists(os.path.join(base_path, settings_path)):
            name = _parse_config_name(os.path.join(base_path, settings_path))
        elif settings_path:
            name = os.path.splitext(os.path.basename(settings_path))[0]
        if knowledge_path and os.path.exists(os.path.join(base_path, knowledge_path)):
            k_dim, k_min, k_max = _parse_knowledge(os.path.join(base_path, knowledge_path))
        if problem is None:
            if name is None:
                raise ValueError("either `problem` or a settings file naming a known configuration is required")
            problem = problems.get_problem(name)
        self._problem = problem
        self._problem_name = problem.name
        self._dimension = problem.dimension if k_dim is None else k_dim
        if self._dimension != problem.dimension:
            raise ValueError("knowledge file dimensionality does not match the problem definition")
        self._min_level = min_level if min_level is not None else (k_min if k_min is not None else problem.min_level)
        self._max_level = max_level if max_level is not None else (k_max if k_max is not None else problem.max_level)
        self._original_min_level, self._original_max_level = self._min_level, self._max_level
        if self._solver_iteration_limit is None:
            self._solver_iteration_limit = problem.solver_iteration_limit
        self._parameters: Dict = {}
        self._device = device
        self._engine_min_level = engine_min_level
        self._resident_inputs = resident_inputs
        # host inputs of evaluation i+1 travel to a device staging buffer on a copy stream while
        # evaluation i solves (one H2D copy per evaluation, off the critical path)
        self._prefetch_inputs = prefetch_inputs
        self._stage = None
        # population.SoloGate shared with the other generators of this process, or None: when set, the
        # solve runs overlapped with theirs (results do not depend on that) and the *time* objective is
        # measured by replaying the captured cycle while the others pause
        self.solo_timing = None
        self.solo_timing_cycles = 3
        self._workspace: Optional[Workspace] = None
        self._workspace_key = None
        self._inputs_key = None
        self._host_inputs = None
        self._front_end = None
        self.last_residual_history: List[float] = []
        self.last_launch_count = 0
        self.last_program = None
        self.last_error = None
        self.last_sweep = []
        self._default_cycle_cache = {}
        self._operator_cache = {}
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._update_grids()

    # -- constructor side products -----------------------------------------------------------
    def _update_grids(self):
        size = 2 ** self._max_level
        h = 1 / (2 ** self._max_level)
        n = self._problem.n_fields
        self._coarsening_factor = [tuple([2] * self._dimension) for _ in range(n)]
        self._finest_grid = [ir.Grid(tuple([size] * self._dimension), tuple([h] * self._dimension), self._max_level)
                             for _ in range(n)]
        self._front_end = None

    @property
    def uses_FAS(self):
        return False

    @property
    def problem(self):
        return self._problem

    @property
    def problem_name(self):
        return self._problem_name

    @property
    def dimension(self):
        return self._dimension

    @property
    def finest_grid(self):
        return self._finest_grid

    @property
    def coarsening_factor(self):
        return self._coarsening_factor

    @property
    def min_level(self):
        return self._min_level

    @property
    def max_level(self):
        return self._max_level

    @property
    def mpi_rank(self):
        return self._mpi_rank

    @property
    def solver_iteration_limit(self):
        return self._solver_iteration_limit

    @property
    def compiler_available(self):
        return True

    def _reference_front_end(self):
        """equations / operators / fields in the form `generate_primitive_set` consumes
        (grammar/multigrid.py:409-478).  They are reference-typed when the reference package
        is importable (drop-in under the unmodified Optimizer), else light records."""
        if self._front_end is None:
            from .front_end import build_front_end
            self._front_end = build_front_end(self._problem, self._min_level, self._max_level, self._parameters)
        return self._front_end

    @property
    def equations(self):
        return self._reference_front_end()[0]

    @property
    def operators(self):
        return self._reference_front_end()[1]

    @property
    def fields(self):
        return self._reference_front_end()[2]

    # -- interface used by the optimizer ----------------------------------------------------------
    def generate_storage(self, min_level: int, max_level: int, finest_grids):
        storage = []
        grids = list(finest_grids)
        top = grids[0].level
        for _ in range(min_level, max_level + 1):
            storage.append(CycleStorage(self._problem, grids, top))
            grids = [ir.Grid(tuple(s // 2 for s in g.size), tuple(h * 2 for h in g.spacing), g.level - 1) for g in grids]
        return storage

    def initialize_code_generation(self, min_level: int, max_level: int, **_ignored):
        """One-time set-up per level range (exastencils.py:445-466): here, allocation of the
        level hierarchy on the GPU.  Extra keyword arguments are accepted because the reference's
        own caller passes one the reference does not take (optimization/program.py:917)."""
        t0 = time.time()
        self._ensure_workspace(min_level, max_level)
        if self._counter == 0:
            self._counter += 1
            self._average_generation_time += (time.time() - t0 - self._average_generation_time) / self._counter
        return None

    def reinitialize(self, min_level, max_level, global_expressions):
        """Level generalisation + PDE parameter update (exastencils.py:196-215)."""
        self._min_level, self._max_level = min_level, max_level
        self._parameters.update(global_expressions or {})
        self._update_grids()
        self._inputs_key = None

    def _device_ordinal(self):
        if self._device is not None:
            return self._device
        return int(os.environ.get("LOCAL_RANK", "0"))

    def _ensure_workspace(self, min_level, max_level):
        lo = min_level if self._engine_min_level is None else min(self._engine_min_level, min_level)
        key = (lo, max_level)
        if self._workspace is None or self._workspace_key != key:
            if self._workspace is not None:
                self._workspace.close()
            p = self._problem
            self._workspace = Workspace(p.dimension, lo, max_level, p.n_fields, p.dtype, self._device_ordinal())
            self._workspace_key = key
            self._inputs_key = None
        return self._workspace

    def _configure(self, ws: Workspace, params):
        p = self._problem
        merged = dict(p.parameters)
        merged.update(params)
        for level in range(ws.min_level, ws.max_level + 1):
            ws.set_operator(level, lowering.operator_table(p.system_operator(level, 0, merged)))
        ws.set_variable_coefficient(p.variable_coefficient_spec(merged))
        ws.set_boundary(p.boundary, merged.get("k", 0.0))
        ws.set_coarse_solver(p.solver.cgs, p.solver.cgs_max_iterations, p.solver.cgs_target_reduction)

    def _load_inputs(self, ws: Workspace, params, rhs=None, u0=None):
        key = (ws.max_level, tuple(sorted((k, complex(v)) for k, v in params.items())), id(rhs), id(u0))
        if self._host_inputs is None or self._host_inputs[0] != key:
            us, fs = self._problem.initial_fields(ws.max_level, params)
            if rhs is not None:
                fs = rhs
            if u0 is not None:
                us = u0
            self._host_inputs = (key, us, fs)
        if self._resident_inputs and self._inputs_key == key:
            ws.restore_initial_guess()
            return
        _, us, fs = self._host_inputs
        staged = self._prefetch_inputs and self._stage is not None and self._stage["key"] == key
        if staged:
            self._stage["event"].synchronize()
        for i in range(ws.n_fields):
            if staged:
                ws.upload_device("rhs", ws.max_level, i, self._stage["rhs"][i].data_ptr())
                ws.upload_device("sol", ws.max_level, i, self._stage["sol"][i].data_ptr())
            else:
                ws.upload("rhs", ws.max_level, i, fs[i])
                ws.upload("sol", ws.max_level, i, us[i])
                self.h2d_bytes += fs[i].nbytes + us[i].nbytes
        ws.save_initial_guess()
        self._inputs_key = key
        if self._prefetch_inputs:
            self._start_prefetch(ws, key, us, fs)

    def _start_prefetch(self, ws: Workspace, key, us, fs):
        """Asynchronous H2D copy of the next evaluation's inputs (the same host buffers, as every
        individual of a population is evaluated on the same problem) into dense staging arrays."""
        import torch
        device = torch.device("cuda", ws.device or 0)
        if self._stage is None or self._stage["shape"] != us[0].shape:
            dtype = torch.from_numpy(us[0][:1]).dtype
            self._stage = {"shape": us[0].shape, "stream": torch.cuda.Stream(device=device),
                           "rhs": [torch.empty(us[0].shape, dtype=dtype, device=device) for _ in range(ws.n_fields)],
                           "sol": [torch.empty(us[0].shape, dtype=dtype, device=device) for _ in range(ws.n_fields)]}
        st = self._stage
        with torch.cuda.stream(st["stream"]):
            for i in range(ws.n_fields):
                st["rhs"][i].copy_(torch.from_numpy(fs[i]), non_blocking=True)
                st["sol"][i].copy_(torch.from_numpy(us[i]), non_blocking=True)
                self.h2d_bytes += fs[i].nbytes + us[i].nbytes
            st["event"] = torch.cuda.Event()
            st["event"].record(st["stream"])
        st["key"] = key

    def set_inputs(self, rhs=None, u0=None):
        """Replace the problem's analytic RHS / initial guess (synthetic throughput inputs)."""
        self._custom_rhs, self._custom_u0 = rhs, u0
        self._inputs_key = None
        self._host_inputs = None

    _custom_rhs = None
    _custom_u0 = None

    def evaluate_program(self, program: lowering.Program, params, infinity=1e100, evaluation_samples=3,
                         max_iterations=None):
        """Solve `evaluation_samples` times; average like `evaluate` (exastencils.py:417-443)."""
        p = self._problem
        if p.outer not in ("mg", "bicgstab"):
            raise EngineError("outer solver '%s' is not executable by the engine" % p.outer)
        ws = self._ensure_workspace(program.min_level, program.max_level)
        if ws.max_level != program.max_level:
            raise EngineError("program finest level differs from the workspace")
        self._configure(ws, params)
        for level, table in program.operators.items():
            if ws._operators.get(level) != table:
                raise EngineError(f"the individual's operator on level {level} differs from the problem's")
        ws.clear_programs()
        default_cycle = None
        if ws.min_level < program.min_level:
            def default_cycle(level, _cache=self._default_cycle_cache):
                # the `.exa3` solver block below the grammar's coarsest level: same for every individual
                key = (p.name, level, ws.min_level, tuple(sorted(params.items())))
                if key not in _cache:
                    _cache[key] = lowering.fuse(lowering.default_cycle_ops(p, level, ws.min_level, params))
                return _cache[key]
        cycle = ws.build_cycle(program.ops, default_cycle)
        self.last_program = program
        # wall-clock cap of one run of the solver (`timeout_evaluate`, exastencils.py:42, 424-431)
        ws.set_time_limit(self.timeout_evaluate)
        tol = p.solver.target_reduction
        max_its = p.solver.max_iterations if max_iterations is None else max_iterations
        total_time = total_rho = total_n = 0.0
        if p.outer == "bicgstab":
            # unshifted operator of the finest level (Helmholtz/2D_FD_Helmholtz_fromL3.exa3:55-62)
            merged = dict(p.parameters)
            merged.update(params)
            ws.set_outer_operator(lowering._canon(stencils.helmholtz_2d(p.spacing(program.max_level), merged["k"], 1.0)))
        gate = self.solo_timing
        shared = gate.shared if gate is not None else contextlib.nullcontext
        for sample in range(evaluation_samples):
            if sample == 0 and p.outer == "mg":
                self._load_inputs(ws, params, self._custom_rhs, self._custom_u0)
            elif p.outer == "mg":
                ws.restore_initial_guess()
            else:
                self._inputs_key = None          # the preconditioner calls overwrite the right-hand side field
                self._load_inputs(ws, params, self._custom_rhs, self._custom_u0)
            if p.outer == "bicgstab":
                # small grids: the whole solve is the timed unit, so it runs alone when a gate is set
                with (gate.exclusive() if gate is not None else contextlib.nullcontext()):
                    history, converged, ms = ws.solve_bicgstab(cycle, tol, max_its)
                self.d2h_bytes += 8 * len(history)
                self.last_residual_history = history
                rho, n = summarise_bicgstab(history, converged, infinity, self._solver_iteration_limit)
            else:
                with shared():
                    history, converged, ms = ws.solve(cycle, tol, max_its)
                self.d2h_bytes += 8 * len(history)
                self.last_residual_history = history
                rho, n = summarise_history(history, len(history) - 1, converged, infinity,
                                           self._solver_iteration_limit)
                if gate is not None and len(history) > 1:
                    with gate.exclusive():
                        ms = ws.run_cycles(cycle, self.solo_timing_cycles) / self.solo_timing_cycles * (len(history) - 1)
            if n >= infinity or rho > 1:
                return ms, rho, n
            if math.isinf(rho) or math.isnan(rho):
                return infinity, infinity, infinity
            total_time += ms
            total_rho += rho
            total_n += n
        self.last_launch_count = ws.launch_count()
        return total_time / evaluation_samples, total_rho / evaluation_samples, total_n / evaluation_samples

    def generate_and_evaluate(self, expression, storages, min_level: int, max_level: int, solver_program: str,
                              infinity=1e100, evaluation_samples=3, global_variable_values=None):
        """exastencils.py:485-537.  A bad individual never raises: constructs the engine cannot execute,
        an evaluation that exceeds `evaluation_timeout` seconds and per-individual engine errors all give
        the sentinel `(infinity,)*3`, like the reference's failed code generation / non-zero exit /
        `subprocess.TimeoutExpired` (:424-431, :493-507).  Only a broken engine (CUDA failure, out of
        memory, missing library: `EngineError.fatal`) propagates."""
        if global_variable_values is None:
            global_variable_values = {}
        mapping = dict(self._parameters)
        mapping.update(global_variable_values)
        n = 3 if "k" in global_variable_values else 1     # :518-521: k, 2k, 4k on the same grid
        avg_t = avg_rho = avg_n = 0.0
        self.last_sweep = []
        for _ in range(n):
            t0 = time.time()
            try:
                # The stencil values of the IR are numbers, so the individual is lowered again for every
                # wave number: operators and local-solve matrices are rebuilt from the problem definition
                # at the current parameter values (the reference substitutes the global variable in the
                # generated program instead, :518-532).
                override = None
                if self._problem.parameters:
                    # PDE parameters are globals of the generated program in the reference, not properties
                    # of the individual: the problem's / generator's / caller's values apply
                    merged = dict(self._problem.parameters)
                    merged.update(mapping)
                    key = tuple(sorted((name, complex(v)) for name, v in merged.items()))

                    def override(level, _key=key, _merged=merged):
                        if (level, _key) not in self._operator_cache:
                            if len(self._operator_cache) > 256:
                                self._operator_cache.clear()
                            self._operator_cache[(level, _key)] = self._problem.system_operator(level, 0, _merged)
                        return self._operator_cache[(level, _key)]
                program = lowering.lower(expression, min_level, max_level, override)
            except lowering.LoweringError:
                return infinity, infinity, infinity
            self._counter += 1
            self._average_generation_time += (time.time() - t0 - self._average_generation_time) / self._counter
            try:
                runtime, rho, its = self.evaluate_program(program, mapping, infinity, evaluation_samples)
            except EngineError as exc:
                if exc.fatal:
                    raise
                self.last_error = str(exc)
                return infinity, infinity, infinity
            self.last_sweep.append((dict(mapping), runtime, rho, its, list(self.last_residual_history)))
            avg_t += runtime
            avg_rho += rho
            avg_n += its
            if its >= infinity or rho > 1:
                return avg_t, avg_rho, avg_n
            if n > 1:
                mapping["k"] *= 2
        return avg_t / n, avg_rho / n, avg_n / n

    def generate_cycle_function(self, expression, storages, min_level: int, level, max_level: int,
                                use_global_weights=False) -> str:
        from .exaslang import emit_cycle_function
        return emit_cycle_function(self._problem, expression, min_level, level, max_level, self._cycle_name,
                                   self._parameters, use_global_weights)

    def close(self):
        if self._stage is not None:
            self._stage["event"].synchronize()
            self._stage = None
        if self._workspace is not None:
            self._workspace.close()
            self._workspace = None
