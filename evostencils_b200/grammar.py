"""Grammar productions: derivation tree (grammar string) -> IR `Cycle` graph.

Host-side mirror of /root/reference/evostencils/grammar/multigrid.py:
  * productions `residual`, `coarsening`, `update_with_coarse_grid_correction`,
    `decoupled_jacobi`, `collective_jacobi`, `collective_block_jacobi`,
    `correct_with_coarse_grid_solver`                       (multigrid.py:253-343)
  * per-level registration and terminal naming             (multigrid.py:238-251, 345-385)
  * block-shape terminals                                   (multigrid.py:388-407)
  * relaxation table `linspace(0.1, 1.9, 37)`               (multigrid.py:428)
  * typed grow generation with a size limit                 (grammar/gp.py:6-52)

The persisted / wire format of a solver is the grammar string
(scripts/optimize.py:177-179, optimization/program.py:918); `Grammar.compile(string)`
evaluates it exactly the way the reference does: `eval(string, pset.context, {})`.
No `deap` is needed: `Grammar` carries its own typed production table.
"""
from __future__ import annotations

import itertools
import random
from functools import reduce
from typing import Dict, List, Optional

import numpy as np

from . import ir
from . import stencils as st
from .problems import ProblemSpec


class Production:
    __slots__ = ("name", "args", "ret", "value")

    def __init__(self, name, args, ret, value=None):
        self.name = name
        self.args = tuple(args)
        self.ret = ret
        self.value = value

    @property
    def arity(self):
        return len(self.args)

    def format(self, *args):
        if self.arity == 0:
            return self.name
        return f"{self.name}({', '.join(args)})"


class Grammar:
    """Typed production table + evaluation context for one problem and level range.

    `depth` = max_level - min_level levels are spanned by productions; the coarse-grid
    solver terminal `CGS_{depth}` lives on `min_level` (multigrid.py:409-478)."""

    def __init__(self, problem: ProblemSpec, min_level: int, max_level: int,
                 maximum_local_system_size: int = 8, relaxation_factor_samples: int = 37,
                 enable_partitioning: bool = True, parameters: Optional[Dict] = None,
                 coarse_grid_solver_expression=None, FAS: bool = False,
                 krylov_methods: tuple = (), krylov_iterations: tuple = (2, 4, 8)):
        """`FAS` selects the productions of the full approximation scheme (multigrid.py:275-299,
        323-341, 360-375): Jacobi-Picard / Jacobi-Newton smoothers, the restricted fine solution
        carried into the coarse right-hand side and subtracted again before the prolongation.
        `krylov_methods` (off by default: the reference's grammar has no production for them) adds
        `<method>_{d}(iterations, C_d) -> S_d` productions that put a KrylovSubspaceMethod
        (ir/krylov_subspace.py:4-45) in the place of a smoother: names from
        {"conjugate_gradient", "bicgstab", "minres", "conjugate_residual"}."""
        depth = max_level - min_level
        assert depth >= 1, "The maximum number of levels must be greater zero"
        self.krylov_methods = tuple(krylov_methods)
        self.problem = problem
        self.min_level = min_level
        self.max_level = max_level
        self.depth = depth
        self.parameters = dict(parameters or {})
        self.relaxation_factor_interval = np.linspace(0.1, 1.9, relaxation_factor_samples)
        self.context: Dict[str, object] = {"__builtins__": None}
        self.primitives: Dict[str, List[Production]] = {}
        self.terminals: Dict[str, List[Production]] = {}
        self.ret = "S_0"
        self.FAS = FAS

        dim = problem.dimension
        nf = problem.n_fields
        fields = problem.field_names
        r_names = problem.restriction_names or [f"gen_restrictionForRes_{f}" for f in fields]
        p_names = problem.prolongation_names or [f"gen_prolongationForSol_{f}" for f in fields]

        finest = problem.grids(max_level)
        self.approximation = ir.Approximation("x", finest, fields)
        self.rhs = ir.RightHandSide("b", finest, problem.rhs_names)
        self._term("u_and_f", "Sg_0", (self.approximation, self.rhs))
        self._term("single", "Partitioning", ir.Single)
        if enable_partitioning:
            self._term("red_black", "Partitioning", ir.RedBlack)
        for i in range(relaxation_factor_samples):
            self._term(repr(i), "RelaxationFactorIndex", i, named=False)
        if self.krylov_methods:
            for k in krylov_iterations:
                self._term(f"its_{k}", "KrylovIterations", int(k))
        if not FAS:
            self._add_block_shapes(nf, dim, maximum_local_system_size)
        else:
            for k in (1, 2, 3, 4):                                        # multigrid.py:446-449
                self._term(repr(k), "NewtonSteps", k, named=False)

        self.operators = []
        for d in range(depth + 1):
            self.operators.append(problem.system_operator(max_level - d, d, self.parameters))
        for d in range(depth):
            fine = problem.grids(max_level - d)
            coarse = problem.grids(max_level - d - 1)
            restriction = ir.Restriction(f"R_{d}", fine, coarse, [st.full_weighting(dim)] * nf, r_names)
            prolongation = ir.Prolongation(f"P_{d + 1}", fine, coarse, [st.multilinear_interpolation(dim)] * nf, p_names)
            coarsest = d == depth - 1
            cgs = None
            if coarsest:
                cgs = ir.CoarseGridSolver("Coarse-Grid Solver", self.operators[d + 1], coarse_grid_solver_expression)
            self._add_level(d, self.operators[d], self.operators[d + 1], restriction, prolongation, cgs, coarse, coarsest)

    # -- table construction --------------------------------------------------------------
    def _term(self, name, type_, value, named=True):
        self.terminals.setdefault(type_, []).append(Production(name, (), type_, value))
        self.primitives.setdefault(type_, [])
        if named:
            assert name not in self.context
            self.context[name] = value

    def _prim(self, name, args, ret, fn):
        assert name not in self.context or self.context[name] is fn
        self.context[name] = fn
        self.primitives.setdefault(ret, []).append(Production(name, args, ret))
        self.terminals.setdefault(ret, [])
        for a in args:
            self.primitives.setdefault(a, [])
            self.terminals.setdefault(a, [])

    def _add_block_shapes(self, n_fields, dimension, maximum_local_system_size):
        per_field = list(itertools.product(range(1, maximum_local_system_size + 1), repeat=dimension))
        for perm in itertools.product(*([per_field] * n_fields)):
            n_terms = sum(reduce(lambda x, y: x * y, shape) for shape in perm)
            if n_fields < n_terms <= maximum_local_system_size:
                self._term(repr(perm), "BlockShape", perm, named=False)
        self.terminals.setdefault("BlockShape", [])
        self.primitives.setdefault("BlockShape", [])

    def _add_level(self, d, operator, coarse_operator, restriction, prolongation, cgs, coarse_grid, coarsest):
        relax = self.relaxation_factor_interval
        scalar_equation = self.problem.n_fields == 1
        S, Sg, C, Cg = f"S_{d}", f"Sg_{d}", f"C_{d}", f"Cg_{d}"
        S2, Sg2, C2, Cg2 = f"S_{d + 1}", f"Sg_{d + 1}", f"C_{d + 1}", f"Cg_{d + 1}"

        if not coarsest:
            self._term(f"zero_{d + 1}", f"x_{d + 1}", ir.ZeroApproximation(coarse_grid))
            self._term(f"A_{d + 1}", f"A_{d + 1}", coarse_operator)
        self._term(prolongation.name, f"P_{d + 1}", prolongation)
        self._term(restriction.name, f"R_{d}", restriction)

        # state transitions (multigrid.py:253-343); cycles are mutated in place
        def residual(state):
            approximation, rhs = state
            return ir.Cycle(approximation, rhs, ir.Residual(operator, approximation, rhs),
                            predecessor=approximation.predecessor)

        def update(relaxation_factor_index, partitioning, cycle):
            cycle.relaxation_factor = relax[relaxation_factor_index]
            cycle.partitioning = partitioning
            return cycle, cycle.rhs

        def smoothing(relaxation_factor_index, partitioning, make_smoother, cycle):
            assert isinstance(cycle.correction, ir.Residual), "Invalid production: expected residual"
            cycle.correction = ir.Multiplication(ir.Inverse(make_smoother(cycle.correction.operator)), cycle.correction)
            return update(relaxation_factor_index, partitioning, cycle)

        def decoupled_jacobi(relaxation_factor_index, partitioning, cycle):
            return smoothing(relaxation_factor_index, partitioning, ir.generate_decoupled_jacobi, cycle)

        def collective_jacobi(relaxation_factor_index, partitioning, cycle):
            return smoothing(relaxation_factor_index, partitioning, ir.generate_collective_jacobi, cycle)

        def collective_block_jacobi(relaxation_factor_index, block_shape, cycle):
            return smoothing(relaxation_factor_index, ir.Single,
                             lambda op: ir.generate_collective_block_jacobi(op, block_shape), cycle)

        def jacobi_picard(relaxation_factor_index, partitioning, cycle):
            return smoothing(relaxation_factor_index, partitioning, ir.generate_jacobi_picard, cycle)

        def jacobi_newton(relaxation_factor_index, partitioning, n_newton_steps, cycle):
            return smoothing(relaxation_factor_index, partitioning,
                             lambda op: ir.generate_jacobi_newton(op, n_newton_steps), cycle)

        def fas_restrict(restriction_operator, cycle):
            # rhs of the coarse problem: R r + A_c (R u)                  (multigrid.py:285-294)
            residual_c = ir.Multiplication(restriction_operator, cycle.correction)
            residual_fas = ir.Multiplication(coarse_operator, ir.Multiplication(restriction_operator, cycle.approximation))
            cycle.correction = ir.Addition(residual_c, residual_fas)
            return cycle

        def coarsening(coarse_operator_, coarse_approximation, restriction_operator, cycle):
            if self.FAS:
                cycle = fas_restrict(restriction_operator, cycle)
            else:
                cycle.correction = ir.Multiplication(restriction_operator, cycle.correction)
            coarse_residual = ir.Residual(coarse_operator_, coarse_approximation, cycle.correction)
            new_cycle = ir.Cycle(coarse_approximation, cycle.correction, coarse_residual)
            new_cycle.predecessor = cycle
            return new_cycle

        def update_with_coarse_grid_correction(relaxation_factor_index, prolongation_operator, state,
                                               restriction_operator=None):
            cycle = state[0]
            if self.FAS:                                                  # multigrid.py:275-283
                correction_fas = ir.Multiplication(restriction_operator, cycle.predecessor.approximation)
                cycle.predecessor.correction = ir.Multiplication(prolongation_operator,
                                                                 ir.Subtraction(cycle, correction_fas))
            else:
                cycle.predecessor.correction = ir.Multiplication(prolongation_operator, cycle)
            return update(relaxation_factor_index, ir.Single, cycle.predecessor)

        def correct_with_coarse_grid_solver(relaxation_factor_index, prolongation_operator, coarse_grid_solver,
                                            restriction_operator, cycle):
            if self.FAS:                                                  # multigrid.py:332-341
                cycle = fas_restrict(restriction_operator, cycle)
                approximation_c = ir.Multiplication(coarse_grid_solver, cycle.correction)
                restricted_solution = ir.Multiplication(restriction_operator, cycle.approximation)
                cycle.correction = ir.Multiplication(prolongation_operator,
                                                     ir.Subtraction(approximation_c, restricted_solution))
                return update(relaxation_factor_index, ir.Single, cycle)
            restricted = ir.Multiplication(restriction_operator, cycle.correction)
            cycle.correction = ir.Multiplication(prolongation_operator,
                                                 ir.Multiplication(coarse_grid_solver, restricted))
            return update(relaxation_factor_index, ir.Single, cycle)

        def both(name, fn, fixed, ins, outs):
            for i, o in zip(ins, outs):
                self._prim(name, fixed + [i], o, fn)

        def krylov(make):
            def production(number_of_iterations, cycle):
                assert isinstance(cycle.correction, ir.Residual), "Invalid production: expected residual"
                cycle.correction = ir.Multiplication(make(cycle.correction.operator, number_of_iterations), cycle.correction)
                return update(len(relax) // 2, ir.Single, cycle)       # the relaxation factor is not applied (exastencils.py:758-768)
            return production

        both(f"residual_{d}", residual, [], [S, Sg], [C, Cg])
        makers = {"conjugate_gradient": ir.generate_conjugate_gradient, "bicgstab": ir.generate_bicgstab,
                  "minres": ir.generate_minres, "conjugate_residual": ir.generate_conjugate_residual}
        for method in self.krylov_methods:
            both(f"{method}_{d}", krylov(makers[method]), ["KrylovIterations"], [C, Cg], [S, Sg])
        if not scalar_equation:
            both(f"decoupled_jacobi_{d}", decoupled_jacobi, ["RelaxationFactorIndex", "Partitioning"], [C, Cg], [S, Sg])
        if not self.FAS:
            both(f"collective_jacobi_{d}", collective_jacobi, ["RelaxationFactorIndex", "Partitioning"], [C, Cg], [S, Sg])
            both(f"collective_block_jacobi_{d}", collective_block_jacobi, ["RelaxationFactorIndex", "BlockShape"], [C, Cg], [S, Sg])
        else:                                                             # multigrid.py:360-364 (argument order as there)
            for i, o in zip([C, Cg], [S, Sg]):
                self._prim(f"jacobi_picard_{d}", ["RelaxationFactorIndex", "Partitioning", i], o, jacobi_picard)
                self._prim(f"jacobi_newton_{d}", ["RelaxationFactorIndex", "Partitioning", "NewtonSteps", i], o, jacobi_newton)
        if not coarsest and self.FAS:                                     # multigrid.py:366-375: the restriction comes last
            for i, o in zip([S2, Sg2], [S, Sg]):
                self._prim(f"update_with_coarse_grid_correction_{d}", ["RelaxationFactorIndex", f"P_{d + 1}", i, f"R_{d}"], o,
                           update_with_coarse_grid_correction)
            both(f"coarsening_{d}", coarsening, [f"A_{d + 1}", f"x_{d + 1}", f"R_{d}"], [C, Cg], [C2, Cg2])
        elif not coarsest:
            both(f"update_with_coarse_grid_correction_{d}", update_with_coarse_grid_correction,
                 ["RelaxationFactorIndex", f"P_{d + 1}"], [S2, Sg2], [S, Sg])
            both(f"coarsening_{d}", coarsening, [f"A_{d + 1}", f"x_{d + 1}", f"R_{d}"], [C, Cg], [C2, Cg2])
        else:
            both(f"correct_with_coarse_grid_solver_{d}", correct_with_coarse_grid_solver,
                 ["RelaxationFactorIndex", f"P_{d + 1}", f"CGC_{d + 1}", f"R_{d}"], [C, Cg], [S, S])
            self._term(f"CGS_{d + 1}", f"CGC_{d + 1}", cgs)

    # -- evaluation ------------------------------------------------------------------------
    def compile(self, grammar_string: str):
        """-> (Cycle, rhs), like `gp.compile` (optimization/program.py:315-317, 918)."""
        return eval(grammar_string, self.context, {})

    # -- typed random generation (grammar/gp.py:6-52) --------------------------------------
    def generate(self, rng: random.Random, min_height=0, max_height=50, type_=None):
        type_ = type_ or self.ret
        expression = []
        height = rng.randint(min_height, max_height)
        stack = [(0, type_)]
        while stack:
            depth, t = stack.pop()
            terminals_available = len(self.terminals[t]) > 0
            if depth < height:
                nodes = self.terminals[t] + self.primitives[t]
            else:
                nodes = self.terminals[t] if terminals_available else self.primitives[t]
            if not nodes:
                raise RuntimeError(f"Neither terminal nor primitive available for {t}")
            choice = rng.choice(nodes)
            for arg in reversed(choice.args):
                stack.append((depth + 1, arg))
            expression.append(choice)
        return expression

    def gen_grow(self, rng: random.Random, min_height=0, max_height=50, size_limit=150):
        result = self.generate(rng, min_height, max_height)
        while len(result) > size_limit:
            result = self.generate(rng, min_height, max_height)
        return result

    @staticmethod
    def to_string(expression) -> str:
        """Prefix list -> grammar string (same layout as deap's PrimitiveTree.__str__)."""
        string = ""
        stack = []
        for node in expression:
            stack.append((node, []))
            while len(stack[-1][1]) == stack[-1][0].arity:
                prim, args = stack.pop()
                string = prim.format(*args)
                if not stack:
                    break
                stack[-1][1].append(string)
        return string


# ----------------------------------------------------------------------------------------
# reference cycles as grammar strings (structure of ir/reference_cycles.py)
# ----------------------------------------------------------------------------------------

def v_cycle_string(depth: int, num_pre: int, num_post: int, relax_index: int, partitioning: str,
                   cgc_relax_index: int = 18, smoother: str = "collective_jacobi") -> str:
    """V(num_pre, num_post) cycle over `depth` grammar levels with a coarse-grid solve.

    Innermost call executes first.  The coarse levels start from the zero approximation
    delivered by `coarsening_d`, which returns a *correction-typed* cycle, so the first
    coarse pre-smoother consumes it directly (multigrid.py:297-299, 381)."""

    def smooth(d, inner):
        return f"{smoother}_{d}({relax_index}, {partitioning}, {inner})"

    # descent
    expr = "u_and_f"
    for d in range(depth):
        if d == 0:
            expr = f"residual_0({expr})"
        for s in range(num_pre):
            expr = smooth(d, expr)
            expr = f"residual_{d}({expr})"
        if d < depth - 1:
            expr = f"coarsening_{d}(A_{d + 1}, zero_{d + 1}, R_{d}, {expr})"
        else:
            expr = f"correct_with_coarse_grid_solver_{d}({cgc_relax_index}, P_{d + 1}, CGS_{d + 1}, R_{d}, {expr})"
    # ascent
    for d in range(depth - 1, -1, -1):
        for s in range(num_post):
            expr = smooth(d, f"residual_{d}({expr})")
        if d > 0:
            expr = f"update_with_coarse_grid_correction_{d - 1}({cgc_relax_index}, P_{d}, {expr})"
    return expr
