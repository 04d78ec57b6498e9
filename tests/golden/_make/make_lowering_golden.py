"""Generate tests/golden/lowering_golden.json from the UNMODIFIED reference front end.

Run in the build container only (needs /root/reference):

    python tests/golden/_make/make_lowering_golden.py

For every case the reference's own `generate_primitive_set` builds the typed primitive
set, the grammar string is evaluated in `pset.context` (what `gp.compile` does,
optimization/program.py:315-317), and `ProgramGenerator.generate_cycle_function`
(code_generation/exastencils.py:318-336) emits the ExaSlang text.  That text is the
known-answer for the lowering pass; it is stored verbatim.  Operator / inter-grid
stencil values, which the reference would parse out of ExaStencils' debug L3 file
(parser.py:25-110; impossible without java), are supplied from
`evostencils_b200.problems` -- only their *names* reach the emitted text.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.abspath(os.path.join(HERE, "..", "..", ".."))
sys.path.insert(0, HERE)                 # deap shim
sys.path.insert(0, "/root/reference")
sys.path.insert(0, REPO)

import sympy  # noqa: E402
from evostencils.ir import base, system  # noqa: E402
from evostencils.grammar import multigrid as mg  # noqa: E402
from evostencils.stencils import constant  # noqa: E402
from evostencils.code_generation.exastencils import ProgramGenerator  # noqa: E402

from evostencils_b200 import problems, stencils as st  # noqa: E402
from evostencils_b200.grammar import v_cycle_string  # noqa: E402


def ref_stencil(s):
    return constant.Stencil(list(s.entries), s.dimension)


def build_reference_front_end(problem, min_level, max_level, maximum_local_system_size):
    dim = problem.dimension
    fields = [sympy.Symbol(n) for n in problem.field_names]
    equations, operators = [], []
    r_names = problem.restriction_names or [f"gen_restrictionForRes_{f}" for f in problem.field_names]
    p_names = problem.prolongation_names or [f"gen_prolongationForSol_{f}" for f in problem.field_names]
    for level in range(min_level, max_level + 1):
        terms = problem.entry_terms(problem.spacing(level), dict(problem.parameters))
        seen = set()
        for i, row in enumerate(terms):
            lhs = []
            for j, entry in enumerate(row):
                for factor, name, sten in entry:
                    f = "" if factor is None else f"{factor!r} * "
                    lhs.append(f"( {f}{name}@{level} * {problem.field_names[j]}@{level} )")
                    if name not in seen:
                        seen.add(name)
                        operators.append(mg.OperatorInfo(name, level, ref_stencil(sten), base.Operator))
            eq = mg.EquationInfo(problem.equation_names[i], level,
                                 "( " + " + ".join(lhs) + f" ) == {problem.rhs_names[i]}@{level}")
            eq._associated_field = fields[i]
            equations.append(eq)
        for rn, pn in zip(r_names, p_names):
            operators.append(mg.OperatorInfo(rn, level, ref_stencil(st.full_weighting(dim)), base.Restriction))
            operators.append(mg.OperatorInfo(pn, level, ref_stencil(st.multilinear_interpolation(dim)), base.Prolongation))
    finest = [base.Grid((2 ** max_level,) * dim, (1 / 2 ** max_level,) * dim, max_level) for _ in fields]
    cf = [(2,) * dim for _ in fields]
    approx = system.Approximation("x", [base.Approximation(f.name, g) for f, g in zip(fields, finest)])
    rhs = system.RightHandSide("b", [base.RightHandSide(n, g) for n, g in zip(problem.rhs_names, finest)])
    pset, terminal_list = mg.generate_primitive_set(approx, rhs, dim, cf, max_level, equations, operators, fields,
                                                    maximum_local_system_size=maximum_local_system_size,
                                                    depth=max_level - min_level)
    pg = object.__new__(ProgramGenerator)   # __init__ shells out to java; only the emitter is wanted
    pg._cycle_name = "gen_mgCycle"
    pg._use_jacobi_prefix = True
    pg._solver_cache = {}
    pg._equations, pg._operators, pg._fields = equations, operators, fields
    pg._dimension = dim
    pg._coarsening_factor = cf
    storages = pg.generate_storage(min_level, max_level, finest)
    return pset, pg, storages


CASES = [
    # (case name, problem, min_level, max_level, max local system size, grammar string)
    ("survey_3_2", "2D_FD_Poisson_fromL2", 2, 4, 4,
     "collective_jacobi_0(18, red_black, residual_0(update_with_coarse_grid_correction_0(9, P_1, "
     "collective_jacobi_1(7, single, residual_1(correct_with_coarse_grid_solver_1(9, P_2, CGS_2, R_1, "
     "residual_1(collective_jacobi_1(7, single, coarsening_0(A_1, zero_1, R_0, residual_0("
     "collective_jacobi_0(18, red_black, residual_0(u_and_f))))))))))))"),
    ("poisson2d_v21_rbgs", "2D_FD_Poisson_fromL2", 5, 9, 4, v_cycle_string(4, 2, 1, 21, "red_black")),
    ("poisson2d_v22_jacobi", "2D_FD_Poisson_fromL2", 5, 9, 4, v_cycle_string(4, 2, 2, 14, "single")),
    ("poisson2d_block22", "2D_FD_Poisson_fromL2", 2, 4, 4,
     "collective_block_jacobi_0(10, ((2, 2),), residual_0(update_with_coarse_grid_correction_0(18, P_1, "
     "collective_block_jacobi_1(12, ((1, 3),), residual_1(correct_with_coarse_grid_solver_1(18, P_2, CGS_2, R_1, "
     "coarsening_0(A_1, zero_1, R_0, residual_0(collective_block_jacobi_0(10, ((4, 1),), residual_0(u_and_f))))))))))"),
    ("poisson2d_wlike", "2D_FD_Poisson_fromL2", 2, 5, 4, None),   # filled in below
    ("poisson3d_v21_rbgs", "3D_FD_Poisson_fromL2", 2, 5, 4, v_cycle_string(3, 2, 1, 23, "red_black")),
    ("elasticity_v21_collective_rb", "2D_FD_LinearElasticity_fromL2", 4, 6, 4,
     v_cycle_string(2, 2, 1, 23, "red_black")),
    ("elasticity_v11_decoupled", "2D_FD_LinearElasticity_fromL2", 4, 6, 4,
     v_cycle_string(2, 1, 1, 14, "single", smoother="decoupled_jacobi")),
    ("elasticity_decoupled_rb", "2D_FD_LinearElasticity_fromL2", 4, 6, 4,
     v_cycle_string(2, 1, 1, 16, "red_black", smoother="decoupled_jacobi")),
    ("elasticity_block", "2D_FD_LinearElasticity_fromL2", 4, 6, 4,
     "collective_block_jacobi_0(12, ((2, 1), (2, 1)), residual_0(update_with_coarse_grid_correction_0(18, P_1, "
     "correct_with_coarse_grid_solver_1(18, P_2, CGS_2, R_1, coarsening_0(A_1, zero_1, R_0, residual_0("
     "collective_block_jacobi_0(12, ((1, 2), (1, 2)), residual_0(u_and_f))))))))"),
    ("helmholtz_v21_rbgs", "2D_FD_Helmholtz_fromL3", 3, 7, 4, v_cycle_string(4, 2, 1, 10, "red_black")),
]


def w_like_string():
    """Level 2 is visited twice from level 1 (W-like shape), built inside-out."""
    e = "residual_0(u_and_f)"
    e = f"residual_0(collective_jacobi_0(21, red_black, {e}))"
    e = f"coarsening_0(A_1, zero_1, R_0, {e})"
    e = f"coarsening_1(A_2, zero_2, R_1, {e})"
    e = f"correct_with_coarse_grid_solver_2(18, P_3, CGS_3, R_2, {e})"
    e = f"collective_jacobi_2(14, single, residual_2({e}))"
    e = f"update_with_coarse_grid_correction_1(18, P_2, {e})"
    e = f"residual_1(collective_jacobi_1(21, red_black, residual_1({e})))"
    e = f"coarsening_1(A_2, zero_2, R_1, {e})"
    e = f"correct_with_coarse_grid_solver_2(18, P_3, CGS_3, R_2, {e})"
    e = f"update_with_coarse_grid_correction_1(18, P_2, {e})"
    e = f"update_with_coarse_grid_correction_0(18, P_1, {e})"
    return f"collective_jacobi_0(21, red_black, residual_0({e}))"


def main():
    out = {}
    for name, pname, lo, hi, mlss, string in CASES:
        if string is None:
            string = w_like_string()
        problem = problems.get_problem(pname)
        pset, pg, storages = build_reference_front_end(problem, lo, hi, mlss)
        expression, _ = eval(string, pset.context, {})
        text = pg.generate_cycle_function(expression, storages, lo, hi, hi)
        out[name] = {"problem": pname, "min_level": lo, "max_level": hi,
                     "maximum_local_system_size": mlss, "grammar": string, "exaslang": text,
                     "terminals": sorted(k for k in pset.context if k != "__builtins__")}
        print(name, len(text.splitlines()), "lines")
    path = os.path.join(HERE, "..", "lowering_golden.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote", os.path.abspath(path))


if __name__ == "__main__":
    main()
