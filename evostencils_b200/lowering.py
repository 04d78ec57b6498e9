"""Lowering: IR `Cycle` graph -> flat op list (the engine's instruction stream).

Follows, rule by rule, the traversal the reference uses to emit ExaSlang
(/root/reference/evostencils/code_generation/exastencils.py:684-925,
`ProgramGenerator.generate_multigrid`), but emits `Op` records instead of text.  The
emission order of that function *defines* the operation sequence of a solver, so it
is reproduced exactly:

  * rhs sub-expressions are emitted once (the reference's `valid` flag,
    exastencils.py:700-703, 746-749, 828-830) -- tracked here in a local id-set, the
    IR is not mutated (the reference resets the flags afterwards, :330-334);
  * `ZeroApproximation` -> zero-fill of the level's solution field (:706-709, 752-755);
  * smoothing = `Inverse(L) * Residual` -> local-solve statements, independent
    equations first, one statement each, then the coupled block (:769-822,
    transformations.py:51-145); `Single` partitioning -> Jacobi (double buffered),
    `RedBlack` -> coloured in-place sweep, colour (sum i_d) % 2 == 0 first (:659-667, :780-790);
  * `R * x` writes the coarse rhs field, `P * x` inside a cycle is `sol += w * P * src`
    (:727-743, :855-873); coarse-grid solver call (:874-911);
  * field naming: solution field on the finest level is the PDE unknown, on coarser
    levels the error field; correction field == error field (:295-316).

The function is duck-typed (class names via the MRO, attribute names) so it accepts
graphs made of `evostencils_b200.ir` objects or of the reference's own
`evostencils.ir.base/system` objects.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

from . import stencils as st

# field kinds of a level's storage (exastencils.py:14-26)
SOL, RHS, RES, COR = "sol", "rhs", "res", "cor"


# Krylov subspace methods the engine executes as fixed-iteration smoothers (ir/krylov_subspace.py:32-45 also
# names MinRes and ConjugateResidual; individuals that use them get the sentinel fitness)
KRYLOV_METHODS = ("ConjugateGradient", "BiCGStab")


class LoweringError(RuntimeError):
    """The graph uses a construct the engine (or ExaStencils) cannot execute."""


@dataclass
class LocalSolve:
    """One `solve locally` statement (exastencils.py:607-633, 769-822).

    `slots`: unknowns of the local system as (field, index-in-tile); `matrix[s][t]` is
    the coefficient of unknown t in the equation of slot s (from the smoothing
    operator L).  The statement visits every tile of shape `period` (anchored at node
    index 0) and sets u_s += omega * (L^-1 r)_s with r = f - A u evaluated at the slots'
    nodes; unknowns on Dirichlet boundary nodes are dropped from the local system."""
    level: int
    omega: float
    jacobi: bool                      # True: read old values (double buffer); False: in place
    colored: bool                     # True: red-black, colour 0 = (sum i_d) % 2 == 0 first
    period: Tuple[int, ...]
    slots: Tuple[Tuple[int, Tuple[int, ...]], ...]
    matrix: Tuple[Tuple[object, ...], ...]
    kind: str = "smooth"

    def key(self):
        return ("smooth", self.level, self.omega, self.jacobi, self.colored, self.period, self.slots)


@dataclass
class Op:
    kind: str                         # residual | restrict | prolong_add | prolong | zero | cgs | copy | richardson | krylov
    level: int                        # level of the field that is written
    omega: float = 1.0
    src: str = ""                     # source field kind
    dst: str = ""                     # destination field kind
    inner: Optional[list] = None      # cgs on a level above the engine's min level: op list of the inner cycle
    method: str = ""                  # krylov: "ConjugateGradient" | "BiCGStab"
    iterations: int = 0               # krylov: number of iterations

    def key(self):
        return (self.kind, self.level, self.omega, self.src, self.dst, self.method, self.iterations)


@dataclass
class Program:
    ops: List[object]
    min_level: int
    max_level: int
    n_fields: int
    dimension: int
    operators: Dict[int, list] = field(default_factory=dict)      # level -> F x F canonical stencil entries

    def signature(self):
        return tuple(op.key() for op in self.ops)


# ----------------------------------------------------------------------------------------
# duck typing helpers
# ----------------------------------------------------------------------------------------

def _is(obj, *names) -> bool:
    return any(c.__name__ in names for c in type(obj).__mro__)


def _grids(expr):
    g = expr.grid
    return list(g) if isinstance(g, (list, tuple)) else [g]


def _level(expr) -> int:
    return _grids(expr)[0].level


def _partitioning_name(p) -> str:
    get = getattr(p, "get_name", None)
    if get is None:
        raise LoweringError(f"unknown partitioning {p!r}")
    name = get()
    if name not in ("single", "red_black"):
        # exastencils.py:668-681 names FourWay/NineWay/... which do not exist in ir/partitioning.py
        raise LoweringError(f"unsupported partitioning {name}")
    return name


def _numeric(v):
    """Stencil values may be sympy numbers when they come from the reference parser."""
    if isinstance(v, (int, float, complex)):
        return v
    try:
        c = complex(v)
    except TypeError as e:  # free symbols left (e.g. an unset PDE parameter)
        raise LoweringError(f"non-numeric stencil value {v!r}") from e
    return c.real if c.imag == 0 else c


def _canon(stencil):
    return tuple((o, _numeric(v)) for o, v in st.canonical_entries(stencil))


def _entry_stencil(entry):
    s = entry.generate_stencil()
    if s is None:
        raise LoweringError(f"operator entry {entry!r} has no stencil")
    # the reference's multiple.add/scale wrap constant stencils into period-1 periodic ones
    if hasattr(s, "constant_stencils") and all(p == 1 for p in st.periodic_shape(s)):
        level = s.constant_stencils
        for _ in range(s.dimension):
            level = level[0]
        s = level
    return s


def operator_table(system_operator) -> list:
    """F x F canonical stencils of a block operator."""
    return [[_canon(_entry_stencil(e)) for e in row] for row in system_operator.entries]


_FW_1D = {-1: 0.25, 0: 0.5, 1: 0.25}
_ML_1D = {-1: 0.5, 0: 1.0, 1: 0.5}


def _check_intergrid(op, weights_1d, what):
    for i, row in enumerate(op.entries):
        entry = row[i]
        sten = entry.generate_stencil()
        if sten is None:
            continue  # no values attached (FAS-style front end): default weights are implied
        got = {tuple(o): float(_numeric(v)) for o, v in sten.entries}
        dim = len(next(iter(got)))
        ref = st.full_weighting(dim) if weights_1d is _FW_1D else st.multilinear_interpolation(dim)
        want = {o: v for o, v in ref.entries}
        if got != want:
            raise LoweringError(f"{what} stencil differs from the engine's tensor-product weights")


# ----------------------------------------------------------------------------------------
# smoother analysis (transformations.py:51-145)
# ----------------------------------------------------------------------------------------

def _periodic_tables(smoothing_operator, system_operator):
    """-> (F x F table of {tile index -> {offset: value}}, per-row period)."""
    rows = system_operator.entries
    nf = len(rows)
    dim = _grids(system_operator)[0].dimension
    one = (1,) * dim
    if _is(smoothing_operator, "ElementwiseDiagonal") or _is(smoothing_operator, "Diagonal"):
        decoupled = not _is(smoothing_operator, "ElementwiseDiagonal")
        tables = []
        for i in range(nf):
            tables.append([])
            for j in range(nf):
                if decoupled and i != j:
                    tables[-1].append({(0,) * dim: {}})
                    continue
                full = _entry_stencil(smoothing_operator.operand.entries[i][j])
                diag = {tuple(o): _numeric(v) for o, v in full.entries if not any(o) and v != 0}
                tables[-1].append({(0,) * dim: diag})
        return tables, [one] * nf
    if hasattr(smoothing_operator, "entries"):
        tables, periods = [], []
        for i in range(nf):
            tables.append([])
            row_period = None
            for j in range(nf):
                sten = _entry_stencil(smoothing_operator.entries[i][j])
                period = st.periodic_shape(sten) or one
                if row_period is None:
                    row_period = period
                elif period != row_period:
                    raise LoweringError("block shapes differ inside one block row")
                table = {}
                if hasattr(sten, "constant_stencils"):
                    def walk(level, prefix):
                        if len(prefix) == dim:
                            table[prefix] = {tuple(o): _numeric(v) for o, v in level.entries if v != 0}
                            return
                        for k, sub in enumerate(level):
                            walk(sub, prefix + (k,))
                    walk(sten.constant_stencils, ())
                else:
                    table[(0,) * dim] = {tuple(o): _numeric(v) for o, v in sten.entries if v != 0}
                tables[-1].append(table)
            periods.append(row_period)
        return tables, periods
    raise LoweringError("Can not extract equations from smoothing operator")


def local_solve_statements(smoothing_operator, system_operator, level, omega, partitioning) -> List[LocalSolve]:
    tables, periods = _periodic_tables(smoothing_operator, system_operator)
    nf = len(tables)
    dim = len(periods[0])
    import itertools
    # equations keyed (field, tile index); unknowns of an equation = "new" symbols
    equations = {}
    for i in range(nf):
        for idx in itertools.product(*(range(p) for p in periods[i])):
            unknowns = {}
            for j in range(nf):
                for off, val in tables[i][j][idx].items():
                    pos = tuple(a + b for a, b in zip(idx, off))
                    unknowns[(j, pos)] = unknowns.get((j, pos), 0) + val
            equations[(i, idx)] = unknowns
    for key, unknowns in equations.items():
        if key not in unknowns or unknowns[key] == 0:
            raise LoweringError(f"local equation {key} has no diagonal entry")
        for u in unknowns:
            if u not in equations:
                raise LoweringError("local system is not square (block shapes of coupled fields differ)")
    # independence test of transformations.py:124-145
    keys = list(equations)
    independent, dependent = [], []
    for a in keys:
        shared = any(u in equations[b] for b in keys if b != a for u in equations[a])
        (dependent if shared else independent).append(a)
    collective = _is(smoothing_operator, "ElementwiseDiagonal")
    if collective:  # exastencils.py:775-777
        dependent = dependent + independent
        independent = []
    name = _partitioning_name(partitioning)
    colored = name == "red_black"
    jacobi = not colored              # `with jacobi` iff Single (exastencils.py:780-790)

    def statement(group):
        period = periods[group[0][0]]
        for g in group:
            if periods[g[0]] != period:
                raise LoweringError("block shapes of coupled fields differ")
        if colored and any(p != 1 for p in period):
            raise LoweringError("coloured block smoothers are not generated by the grammar")
        matrix = tuple(tuple(equations[s].get(t, 0.0) for t in group) for s in group)
        return LocalSolve(level=level, omega=float(omega), jacobi=jacobi, colored=colored,
                          period=tuple(period), slots=tuple(group), matrix=matrix)

    out = [statement([k]) for k in independent]
    if dependent:
        out.append(statement(dependent))
    return out


# ----------------------------------------------------------------------------------------
# main traversal
# ----------------------------------------------------------------------------------------
class _Lowerer:
    def __init__(self, min_level, max_level, operator_override=None):
        self.min_level = min_level
        self.max_level = max_level
        # level -> SystemOperator that replaces the graph's own system operator on that level: the same
        # individual evaluated for other PDE parameters (k, 2k, 4k of exastencils.py:518-532 -- there
        # the parameter is a global variable of the generated program, here the stencil values are numbers)
        self.operator_override = operator_override
        self.ops: List[object] = []
        self.emitted = set()
        self.operators: Dict[int, list] = {}
        self.n_fields = None
        self.dimension = None

    def sol_kind(self, level):
        return SOL

    def system_operator(self, system_operator, level):
        return self.operator_override(level) if self.operator_override else system_operator

    def smoothing_operator(self, smoothing_operator, level):
        """The smoother marker of the graph, rebuilt around the overriding system operator."""
        if not self.operator_override:
            return smoothing_operator
        from . import ir
        operator = self.operator_override(level)
        if _is(smoothing_operator, "ElementwiseDiagonal"):
            return ir.ElementwiseDiagonal(operator)
        if _is(smoothing_operator, "Diagonal"):
            return ir.Diagonal(operator)
        if hasattr(smoothing_operator, "block_sizes"):
            return ir.BlockDiagonalOperator(operator, smoothing_operator.block_sizes)
        if hasattr(smoothing_operator, "entries"):
            # the reference's block-diagonal operator (ir/smoother.py:13-22) does not record its block
            # shapes: they are the periods of the diagonal entries' stencils
            dim = _grids(operator)[0].dimension
            sizes = tuple(st.periodic_shape(_entry_stencil(row[i])) or (1,) * dim
                          for i, row in enumerate(smoothing_operator.entries))
            return ir.BlockDiagonalOperator(operator, sizes)
        raise LoweringError("the smoothing operator can not be rebuilt for other PDE parameters")

    def note_operator(self, system_operator, level):
        table = operator_table(self.system_operator(system_operator, level))
        if level in self.operators and self.operators[level] != table:
            raise LoweringError(f"two different system operators on level {level}")
        self.operators[level] = table
        self.n_fields = len(table)
        self.dimension = _grids(system_operator)[0].dimension

    def source_field(self, expression) -> str:
        # exastencils.py:594-604
        if _is(expression, "Cycle") or (_is(expression, "Approximation") and not _is(expression, "RightHandSide")):
            return SOL
        if _is(expression, "Residual"):
            return RES
        if _is(expression, "RightHandSide"):
            return RHS
        return COR

    def preamble(self, owner_rhs, rhs_for_type_check, approximation, zero_of):
        if not _is(rhs_for_type_check, "RightHandSide") and id(owner_rhs) not in self.emitted:
            self.emit(owner_rhs)
            self.emitted.add(id(owner_rhs))
        if not _is(approximation, "Approximation"):
            self.emit(approximation)
        if _is(zero_of, "ZeroApproximation"):
            self.ops.append(Op("zero", _level(zero_of), dst=SOL))

    def emit(self, expression):
        if _is(expression, "Cycle"):
            weight = expression.relaxation_factor
            correction = expression.correction
            level = _level(expression)
            if _is(correction, "Residual"):
                self.preamble(expression.rhs, correction.rhs, correction.approximation, expression.approximation)
                self.note_operator(correction.operator, level)
                self.ops.append(Op("richardson", level, omega=float(weight), dst=SOL))
            elif _is(correction, "Multiplication"):
                op1, op2 = correction.operand1, correction.operand2
                if _is(op1, "InterGridOperator"):
                    self.emit(op2)
                    if not _is(op1, "Prolongation"):
                        raise LoweringError("restriction inside a cycle update is not generated by the grammar")
                    _check_intergrid(op1, _ML_1D, "prolongation")
                    self.ops.append(Op("prolong_add", level, omega=float(weight), src=self.source_field(op2), dst=SOL))
                elif _is(op1, "Inverse"):
                    residual = op2
                    self.preamble(residual.rhs, residual.rhs, residual.approximation, expression.approximation)
                    self.note_operator(residual.operator, level)
                    self.ops.extend(local_solve_statements(self.smoothing_operator(op1.operand, level),
                                                           self.system_operator(residual.operator, level), level,
                                                           weight, expression.partitioning))
                elif _is(op1, "KrylovSubspaceMethod"):
                    # exastencils.py:744-768: same preamble as a smoother, then `<Name>_<n>@level()`; the
                    # relaxation factor of the cycle is not applied (the solver updates the solution itself)
                    residual = op2
                    self.preamble(residual.rhs, residual.rhs, residual.approximation, expression.approximation)
                    self.note_operator(residual.operator, level)
                    if op1.name not in KRYLOV_METHODS:
                        raise LoweringError(f"Krylov subspace method '{op1.name}' is not executable by the engine")
                    n = int(op1.number_of_iterations)
                    if n < 1:
                        raise LoweringError("Krylov subspace method with no iterations")
                    self.ops.append(Op("krylov", level, src=RHS, dst=SOL, method=op1.name, iterations=n))
                else:
                    raise LoweringError("Unsupported operator")
            else:
                raise LoweringError("Expected multiplication")
        elif _is(expression, "Residual"):
            self.preamble(expression.rhs, expression.rhs, expression.approximation, expression.approximation)
            level = _level(expression)
            self.note_operator(expression.operator, level)
            self.ops.append(Op("residual", level, dst=RES))
        elif _is(expression, "Multiplication"):
            op1, op2 = expression.operand1, expression.operand2
            if _is(op1, "InterGridOperator"):
                self.emit(op2)
                src = self.source_field(op2)
                if _is(op1, "Restriction"):
                    _check_intergrid(op1, _FW_1D, "restriction")
                    self.ops.append(Op("restrict", _level(expression), src=src, dst=RHS))
                else:
                    _check_intergrid(op1, _ML_1D, "prolongation")
                    self.ops.append(Op("prolong", _level(expression), src=src, dst=COR))
            elif _is(op1, "CoarseGridSolver"):
                self.emit(op2)
                level = min(g.level for g in _grids(op2))
                self.note_operator(op1.operator, level)
                inner = None
                if level != self.min_level:
                    sub = getattr(op1, "expression", None)
                    if sub is not None:
                        # previously optimised cycle for the levels below
                        # (optimization/program.py:863-897: `best_expression`)
                        lw = _Lowerer(self.min_level, level, self.operator_override)
                        lw.emit(sub)
                        inner = lw.ops
                        for lv, table in lw.operators.items():
                            self.operators.setdefault(lv, table)
                    else:
                        inner = []      # engine inserts the problem's default cycle
                self.ops.append(Op("cgs", level, src=RHS, dst=SOL, inner=inner))
            else:
                raise LoweringError("Not implemented")
        else:
            raise LoweringError(f"Not implemented: {type(expression).__name__}")


def lower(expression, min_level: int, max_level: int, operator_override=None) -> Program:
    """IR graph of one solver -> `Program`.  `min_level` is the level of the coarse-grid
    solver, `max_level` the finest level (same meaning as the arguments of
    `generate_cycle_function`, exastencils.py:318-336).  `operator_override(level)` -> system operator
    used instead of the graph's own (same individual, other PDE parameter values)."""
    lw = _Lowerer(min_level, max_level, operator_override)
    lw.emit(expression)
    if lw.n_fields is None:
        raise LoweringError("empty program")
    return Program(lw.ops, min_level, max_level, lw.n_fields, lw.dimension, lw.operators)


# ----------------------------------------------------------------------------------------
# default (reference-solver) cycle as ops
# ----------------------------------------------------------------------------------------

def default_cycle_ops(problem, level: int, min_level: int, parameters=None) -> List[object]:
    """V(numPre, numPost) cycle ExaStencils generates from the `.exa3` solver block
    (e.g. Poisson/2D_FD_Poisson_fromL2.exa3:2-15): coloured point smoother with
    `damping`, undamped coarse-grid correction, CG on `min_level`.  Written as the same
    op vocabulary `lower()` produces (structure: Helmholtz/...exa3:240-272)."""
    from . import ir  # local import: only the markers are needed
    s = problem.solver
    ops: List[object] = []
    partitioning = ir.RedBlack if s.coloring == "red_black" else ir.Single

    def smooth(lv):
        operator = problem.system_operator(lv, 0, parameters)
        return local_solve_statements(ir.ElementwiseDiagonal(operator), operator, lv, s.damping, partitioning)

    def visit(lv):
        if lv == min_level:
            ops.append(Op("cgs", lv, src=RHS, dst=SOL))
            return
        for _ in range(s.num_pre):
            ops.extend(smooth(lv))
        ops.append(Op("residual", lv, dst=RES))
        ops.append(Op("restrict", lv - 1, src=RES, dst=RHS))
        ops.append(Op("zero", lv - 1, dst=SOL))
        visit(lv - 1)
        ops.append(Op("prolong_add", lv, omega=1.0, src=SOL, dst=SOL))
        for _ in range(s.num_post):
            ops.extend(smooth(lv))

    visit(level)
    return ops


# ----------------------------------------------------------------------------------------
# peephole fusion -- used by the CUDA executor, never by the oracle
# ----------------------------------------------------------------------------------------

def fuse(ops: Sequence[object]) -> List[object]:
    """residual -> restrict  ==> residual_restrict (the residual field is still written
    only if a later op reads it); everything else passes through."""
    out: List[object] = []
    i = 0
    n = len(ops)
    while i < n:
        op = ops[i]
        nxt = ops[i + 1] if i + 1 < n else None
        if (isinstance(op, Op) and op.kind == "residual" and isinstance(nxt, Op) and nxt.kind == "restrict"
                and nxt.src == RES and nxt.level == op.level - 1):
            out.append(Op("residual_restrict", nxt.level, src=SOL, dst=RHS))
            i += 2
            continue
        out.append(op)
        i += 1
    return out


def estimate_cycle_bytes(ops: Sequence[object], dimension: int, bytes_per_value: int = 8, default_cycle=None) -> float:
    """Algorithmic HBM bytes of one cycle of `ops` (every input field read once,
    every output written once; interior nodes only; a zero fill that the next smoother of the level
    absorbs is free).  `default_cycle(level)` -> op list the engine runs for a coarse-grid-solver op
    above its min level (the `.exa3` cycle below the grammar's coarsest level).  Static information
    about an individual: orders a population longest-first, and is the numerator of the whole-step
    roofline fraction bench.py prints."""
    total = 0.0
    fused = fuse(ops)
    for i, op in enumerate(fused):
        fine_level = op.level
        if isinstance(op, LocalSolve):
            fields = 3.0                                       # u, f in, u out
        elif op.kind in ("residual", "richardson"):
            fields = 3.0
        elif op.kind in ("residual_restrict", "restrict"):
            fields, fine_level = 2.0 + 0.5 ** dimension, op.level + 1
        elif op.kind in ("prolong_add", "prolong"):
            fields = 2.0 + 0.5 ** dimension
        elif op.kind in ("zero",):
            nxt = fused[i + 1] if i + 1 < len(fused) else None
            fields = 0.0 if isinstance(nxt, LocalSolve) and nxt.level == op.level else 1.0
        elif op.kind == "copy":
            fields = 2.0
        elif op.kind == "cgs":
            inner = op.inner if op.inner else (default_cycle(op.level) if default_cycle else None)
            if inner:
                total += estimate_cycle_bytes(inner, dimension, bytes_per_value, default_cycle)
            continue
        elif op.kind == "krylov":
            # residual (3) + per CG iteration: A p (2), p.Ap (2), x and r updates (6), r.r (1), p update (3); + x += e (3)
            fields = 6.0 + (14.0 if op.method == "ConjugateGradient" else 30.0) * op.iterations
        else:
            fields = 3.0
        total += fields * bytes_per_value * float((2 ** fine_level - 1) ** dimension)
    return total
