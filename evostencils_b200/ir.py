"""Expression IR consumed by the B200 multigrid engine.

This is the *input format* of the engine: the `Cycle` / `Residual` / `Inverse` /
inter-grid graphs that EvoStencils' grammar productions build.  The class and
attribute names mirror the reference so that `lowering.lower()` is duck-typed over both
this module and the reference's own objects:

  * /root/reference/evostencils/ir/base.py:9-724   (Expression, Grid, Operator, Cycle, ...)
  * /root/reference/evostencils/ir/system.py:5-158  (block/system wrappers)
  * /root/reference/evostencils/ir/smoother.py:5-38 (smoother constructors)
  * /root/reference/evostencils/ir/partitioning.py:18-47

Unlike the reference (two modules, `base` for one field and `system` for blocks of
fields) there is a single level here: every approximation / right-hand side / operator
is a *system* object over `n_fields >= 1` fields, which is what the grammar layer always
produces (grammar/multigrid.py:253-343).  Per-field operator entries are small
expression trees over named stencil operators (`Operator`, `Scaling`, `Addition`,
`Multiplication`), evaluated numerically by `generate_stencil()`.
"""
from __future__ import annotations

from typing import List, Sequence

from . import stencils as st


class Expression:
    """Scratch attributes the optimizer / generator mutate (base.py:9-13)."""

    def __init__(self):
        self.lfa_symbol = None
        self.valid = False
        self.runtime = None

    def mutate(self, f, *args):  # default: leaf
        pass


# ----------------------------------------------------------------------------------------
# grids
# ----------------------------------------------------------------------------------------
class Grid:
    """`size` = cells per dimension (2**level), `spacing` = h (base.py:168-196)."""

    def __init__(self, size, spacing, level):
        assert len(size) == len(spacing)
        self.size = tuple(size)
        self.spacing = tuple(spacing)
        self.level = level

    @property
    def dimension(self):
        return len(self.size)

    def __eq__(self, other):
        return isinstance(other, Grid) and self.size == other.size and self.spacing == other.spacing

    def __hash__(self):
        return hash((self.size, self.spacing))

    def __repr__(self):
        return f"Grid({self.size!r}, {self.spacing!r}, {self.level})"


def get_coarse_grid(grids: Sequence[Grid], coarsening_factors) -> List[Grid]:
    """system.py:141-142 / base.py:700-704."""
    out = []
    for g, cf in zip(grids, coarsening_factors):
        out.append(Grid(tuple(s // f for s, f in zip(g.size, cf)),
                        tuple(h * f for h, f in zip(g.spacing, cf)), g.level - 1))
    return out


# ----------------------------------------------------------------------------------------
# per-field operator entries
# ----------------------------------------------------------------------------------------
class Operator(Expression):
    """Named single-field stencil operator (base.py:122-145)."""

    def __init__(self, name, grid: Grid, stencil_generator=None):
        super().__init__()
        self.name = name
        self.grid = grid
        self.stencil_generator = stencil_generator

    def generate_stencil(self):
        if self.stencil_generator is None:
            return None
        return self.stencil_generator.generate_stencil(self.grid)

    def __repr__(self):
        return f"Operator({self.name!r}@{self.grid.level})"


class ConstantStencilGenerator:
    """base.py:719-724."""

    def __init__(self, stencil):
        self._stencil = stencil

    def generate_stencil(self, _grid=None):
        return self._stencil


class Identity(Operator):
    def __init__(self, grid, name="I"):
        super().__init__(name, grid, ConstantStencilGenerator(st.unit_stencil(grid.dimension)))


class ZeroOperator(Operator):
    def __init__(self, grid, name="0"):
        super().__init__(name, grid, ConstantStencilGenerator(st.null_stencil(grid.dimension)))


class Scaling(Expression):
    """factor * operand (base.py:408-449)."""

    def __init__(self, factor, operand):
        super().__init__()
        self.factor = factor
        self.operand = operand

    @property
    def grid(self):
        return self.operand.grid

    def generate_stencil(self):
        return self.operand.generate_stencil().scaled(self.factor)


class Addition(Expression):
    def __init__(self, operand1, operand2):
        super().__init__()
        self.operand1 = operand1
        self.operand2 = operand2

    @property
    def grid(self):
        return self.operand1.grid

    def generate_stencil(self):
        return self.operand1.generate_stencil() + self.operand2.generate_stencil()


class Subtraction(Expression):
    """operand1 - operand2 (base.py `sub`); in FAS cycles: coarse solution minus the restricted fine
    solution (grammar/multigrid.py:276-281, 337-341)."""

    def __init__(self, operand1, operand2):
        super().__init__()
        self.operand1 = operand1
        self.operand2 = operand2

    @property
    def grid(self):
        return self.operand1.grid


class Multiplication(Expression):
    """Binary product node (base.py:385-405).  Used both for operator entries
    (stencil * stencil) and for the cycle-level products `R * residual`,
    `P * cycle`, `Inverse(L) * residual`, `CGS * (...)`."""

    def __init__(self, operand1, operand2):
        super().__init__()
        self.operand1 = operand1
        self.operand2 = operand2

    @property
    def grid(self):
        return self.operand1.grid

    def generate_stencil(self):
        return self.operand1.generate_stencil() @ self.operand2.generate_stencil()

    def mutate(self, f, *args):
        f(self.operand1, *args)
        f(self.operand2, *args)


# ----------------------------------------------------------------------------------------
# system-level entities
# ----------------------------------------------------------------------------------------
class SystemOperator(Expression):
    """n_fields x n_fields block operator (system.py:32-50)."""

    def __init__(self, name, entries):
        super().__init__()
        self.name = name
        self.entries = [list(row) for row in entries]

    @property
    def grid(self):
        return [e.grid for e in self.entries[0]]

    def __repr__(self):
        return f"SystemOperator({self.name!r})"


class Approximation(Expression):
    """List of per-field unknowns on one level (system.py:72-92)."""

    def __init__(self, name, grids: Sequence[Grid], field_names=None):
        super().__init__()
        self.name = name
        self._grid = list(grids)
        self.field_names = list(field_names) if field_names is not None else None

    @property
    def grid(self):
        return self._grid

    @property
    def predecessor(self):
        return None


class RightHandSide(Approximation):
    pass


class ZeroApproximation(Approximation):
    def __init__(self, grids, name="0"):
        super().__init__(name, grids)


class InterGridOperator(SystemOperator):
    """Block-diagonal inter-grid transfer, one stencil per field (system.py:106-120)."""

    def __init__(self, name, fine_grid, coarse_grid, stencils, entry_names=None):
        n = len(fine_grid)
        entry_names = entry_names or [name] * n
        entries = [[_InterGridEntry(entry_names[i], fine_grid[i], coarse_grid[i], stencils[i], type(self).__name__)
                    if i == j else ZeroOperator(fine_grid[i]) for j in range(n)] for i in range(n)]
        super().__init__(name, entries)
        self.fine_grid = list(fine_grid)
        self.coarse_grid = list(coarse_grid)


class _InterGridEntry(Operator):
    def __init__(self, name, fine_grid, coarse_grid, stencil, kind):
        super().__init__(name, coarse_grid if kind == "Restriction" else fine_grid, ConstantStencilGenerator(stencil))
        self.fine_grid = fine_grid
        self.coarse_grid = coarse_grid
        self.kind = kind


class Restriction(InterGridOperator):
    @property
    def grid(self):
        return self.coarse_grid


class Prolongation(InterGridOperator):
    @property
    def grid(self):
        return self.fine_grid


class CoarseGridSolver(Expression):
    """base.py:572-595.  `expression` optionally holds an already-optimised cycle for
    the levels below (multi-run composition, optimization/program.py:863-897)."""

    def __init__(self, name, operator, expression=None):
        super().__init__()
        self.name = name
        self.operator = operator
        self.expression = expression

    @property
    def grid(self):
        return self.operator.grid


# smoother markers (system.py:123-130, smoother.py:5-22)
class Inverse(Expression):
    def __init__(self, operand):
        super().__init__()
        self.operand = operand

    @property
    def grid(self):
        return self.operand.grid


class Diagonal(Expression):
    """Decoupled (per-field) Jacobi: diag of the diagonal blocks only."""

    def __init__(self, operand):
        super().__init__()
        self.operand = operand

    @property
    def grid(self):
        return self.operand.grid


class ElementwiseDiagonal(Diagonal):
    """Collective Jacobi: the n_fields x n_fields point block."""

    def __str__(self):
        return "D"


class BlockDiagonalOperator(SystemOperator):
    """Collective block Jacobi: every entry filtered to non-overlapping tiles
    (smoother.py:13-22).  `block_sizes[i]` is the tile shape of block row i."""

    def __init__(self, operator: SystemOperator, block_sizes):
        self.block_sizes = tuple(tuple(b) for b in block_sizes)
        entries = []
        for i, row in enumerate(operator.entries):
            entries.append([])
            for j, entry in enumerate(row):
                periodic = st.block_diagonal(entry.generate_stencil(), self.block_sizes[i])
                entries[-1].append(Operator(f"{operator.name}_{i}{j}_block_diag", entry.grid,
                                            ConstantStencilGenerator(periodic)))
        super().__init__(f"{operator.name}_block_diag", entries)


class Jacobian(Expression):
    """Derivative of the nonlinear term, used with `n_newton_steps` local Newton updates (system.Jacobian;
    smoother.py:45-46)."""

    def __init__(self, operand, n_newton_steps=1):
        super().__init__()
        self.operand = operand
        self.n_newton_steps = n_newton_steps

    @property
    def grid(self):
        return self.operand.grid


def generate_jacobi_picard(operator):
    """smoother.py:41-42"""
    return ElementwiseDiagonal(operator)


def generate_jacobi_newton(operator, n_newton_steps):
    """smoother.py:45-46"""
    return Addition(ElementwiseDiagonal(operator), Jacobian(operator, n_newton_steps))


def generate_decoupled_jacobi(operator):
    return Diagonal(operator)


def generate_collective_jacobi(operator):
    return ElementwiseDiagonal(operator)


def generate_collective_block_jacobi(operator, block_sizes):
    return BlockDiagonalOperator(operator, block_sizes)


# partitionings (partitioning.py:18-47); the grammar passes the *classes* around
class KrylovSubspaceMethod(Expression):
    """`name` in {ConjugateGradient, BiCGStab, MinRes, ConjugateResidual} applied for a fixed number of
    iterations in place of a smoother: Cycle.correction = Multiplication(KrylovSubspaceMethod, Residual)
    (ir/krylov_subspace.py:4-45; emitted at code_generation/exastencils.py:758-768)."""

    def __init__(self, name, operator, number_of_iterations):
        super().__init__()
        self.name = name
        self.operator = operator
        self.number_of_iterations = number_of_iterations
        self.shape = getattr(operator, "shape", None)

    @staticmethod
    def generate_stencil():
        return None

    @property
    def grid(self):
        return self.operator.grid

    def __repr__(self):
        return f"KrylovSubspaceMethod({self.name!r}, {self.operator!r}, {self.number_of_iterations!r})"


def generate_conjugate_gradient(operator, number_of_iterations):
    return KrylovSubspaceMethod("ConjugateGradient", operator, number_of_iterations)


def generate_bicgstab(operator, number_of_iterations):
    return KrylovSubspaceMethod("BiCGStab", operator, number_of_iterations)


def generate_minres(operator, number_of_iterations):
    return KrylovSubspaceMethod("MinRes", operator, number_of_iterations)


def generate_conjugate_residual(operator, number_of_iterations):
    return KrylovSubspaceMethod("ConjugateResidual", operator, number_of_iterations)


class Single:
    @staticmethod
    def get_name():
        return "single"


class RedBlack:
    @staticmethod
    def get_name():
        return "red_black"


# ----------------------------------------------------------------------------------------
# residual / cycle
# ----------------------------------------------------------------------------------------
class Residual(Expression):
    """rhs - operator * approximation (base.py:598-648)."""

    def __init__(self, operator, approximation, rhs):
        super().__init__()
        self.operator = operator
        self.approximation = approximation
        self.rhs = rhs

    @property
    def grid(self):
        return self.rhs.grid

    def mutate(self, f, *args):
        f(self.rhs, *args)
        f(self.approximation, *args)


class Cycle(Expression):
    """approximation + relaxation_factor * correction (base.py:651-697)."""

    def __init__(self, approximation, rhs, correction=None, partitioning=Single,
                 relaxation_factor=1.0, predecessor=None):
        super().__init__()
        self.approximation = approximation
        self.rhs = rhs
        self.correction = correction
        self.relaxation_factor = relaxation_factor
        self.partitioning = partitioning
        self.predecessor = predecessor
        self.global_id = None

    @property
    def grid(self):
        return self.approximation.grid

    def mutate(self, f, *args):
        f(self.correction, *args)
