"""Generate tests/golden/fas_golden.json from the UNMODIFIED reference FAS front end.

Run in the build container only (needs /root/reference):

    python tests/golden/_make/make_fas_golden.py

The reference's `generate_primitive_set(..., FAS=True)` (grammar/multigrid.py:411-477) builds the typed
primitive set of the nonlinear example, FAS grammar strings are evaluated in `pset.context` (what
`gp.compile` does) and `ProgramGeneratorFAS.generate_cycle_function`
(code_generation/exastencils_FAS.py:428-437) emits the Layer-4 text of the cycle function.  That
text is the known answer for the statement order of `evostencils_b200.fas.v_cycle_ops`.
`ProgramGeneratorFAS.__init__` reads `example_problems/FAS_2D_Basic/*` relative to the working
directory, so the script runs from /root/reference (read-only; nothing is written there)."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.abspath(os.path.join(HERE, "..", "..", ".."))
sys.path.insert(0, HERE)                 # deap shim
sys.path.insert(0, "/root/reference")
sys.path.insert(0, REPO)

os.chdir("/root/reference")

from evostencils.ir import base, system  # noqa: E402
from evostencils.grammar import multigrid as mg  # noqa: E402
from evostencils.code_generation.exastencils_FAS import ProgramGeneratorFAS  # noqa: E402


def fas_v_cycle_string(depth, num_pre, num_post, relax_index, newton_steps=1, partitioning="single", cgc_relax_index=18):
    """V(num_pre, num_post) in the FAS grammar: jacobi_newton smoothers, the restriction operator as an
    extra argument of update_with_coarse_grid_correction (grammar/multigrid.py:300-302, 364-372)."""
    def smooth(d, inner):
        return f"jacobi_newton_{d}({relax_index}, {partitioning}, {newton_steps}, {inner})"

    expr = "u_and_f"
    for d in range(depth):
        if d == 0:
            expr = f"residual_0({expr})"
        for _ in range(num_pre):
            expr = smooth(d, expr)
            expr = f"residual_{d}({expr})"
        if d < depth - 1:
            expr = f"coarsening_{d}(A_{d + 1}, zero_{d + 1}, R_{d}, {expr})"
        else:
            expr = f"correct_with_coarse_grid_solver_{d}({cgc_relax_index}, P_{d + 1}, CGS_{d + 1}, R_{d}, {expr})"
    for d in range(depth - 1, -1, -1):
        for _ in range(num_post):
            expr = smooth(d, f"residual_{d}({expr})")
        if d > 0:
            expr = f"update_with_coarse_grid_correction_{d - 1}({cgc_relax_index}, P_{d}, {expr}, R_{d - 1})"
    return expr


def fas_w_like_string():
    """Level 2 of the grammar is visited twice from level 1; Picard and Newton smoothers, both partitionings."""
    e = "residual_0(u_and_f)"
    e = f"residual_0(jacobi_newton_0(21, red_black, 2, {e}))"
    e = f"coarsening_0(A_1, zero_1, R_0, {e})"
    e = f"coarsening_1(A_2, zero_2, R_1, {e})"
    e = f"correct_with_coarse_grid_solver_2(18, P_3, CGS_3, R_2, {e})"
    e = f"jacobi_picard_2(14, single, residual_2({e}))"
    e = f"update_with_coarse_grid_correction_1(16, P_2, {e}, R_1)"
    e = f"residual_1(jacobi_newton_1(12, single, 3, residual_1({e})))"
    e = f"coarsening_1(A_2, zero_2, R_1, {e})"
    e = f"correct_with_coarse_grid_solver_2(18, P_3, CGS_3, R_2, {e})"
    e = f"update_with_coarse_grid_correction_1(18, P_2, {e}, R_1)"
    e = f"update_with_coarse_grid_correction_0(20, P_1, {e}, R_0)"
    return f"jacobi_picard_0(22, red_black, residual_0(jacobi_newton_0(9, single, 1, residual_0({e}))))"


INLINE_CASES = [
    # (name, depth, grammar string): smoothers emitted inline (fct_smoother=None)
    ("fas_inline_picard_newton", 1,
     "jacobi_picard_0(20, red_black, residual_0(correct_with_coarse_grid_solver_0(18, P_1, CGS_1, R_0, "
     "residual_0(jacobi_newton_0(14, single, 2, residual_0(u_and_f))))))"),
    ("fas_inline_v_two_coarsenings", 2,
     "jacobi_picard_0(20, red_black, residual_0(update_with_coarse_grid_correction_0(18, P_1, "
     "correct_with_coarse_grid_solver_1(18, P_2, CGS_2, R_1, residual_1(jacobi_newton_1(14, single, 2, "
     "coarsening_0(A_1, zero_1, R_0, residual_0(jacobi_newton_0(14, single, 1, residual_0(jacobi_newton_0(14, single, 1, "
     "residual_0(u_and_f))))))))), R_0)))"),
    ("fas_inline_w_like", 3, None),
]


def front_end(pg, depth):
    hi = pg.max_level
    approx = system.Approximation("x", [base.Approximation("u", pg.finest_grid[0])])
    rhs = system.RightHandSide("b", [base.RightHandSide("RHS_u", pg.finest_grid[0])])
    pset, _ = mg.generate_primitive_set(approx, rhs, pg.dimension, pg.coarsening_factor, hi, pg.equations,
                                        pg.operators, pg.fields, depth=depth, FAS=True)
    return pset


def main():
    # fct_smoother = "Smoother": the smoother is the function of the template (.exa4:65-75), as in the
    # generated sample FAS_2D_Basic.exa4:210-269
    pg = ProgramGeneratorFAS("FAS_2D_Basic", "Solution", "RHS", "Residual", "Approximation", "RestrictionNode",
                             "CorrectionNode", "Laplace", "gamSten", "gen_mgCycle", "CGS", "Smoother")
    print("knowledge levels", pg.min_level, pg.max_level, "dimension", pg.dimension)
    out = {}
    for name, depth, pre, post, cgc in (("fas_v22_two_coarsenings", 2, 2, 2, 18), ("fas_v21_three_coarsenings", 3, 2, 1, 18),
                                        ("fas_v11_one_coarsening", 1, 1, 1, 18), ("fas_v12_damped_correction", 2, 1, 2, 12)):
        lo, hi = pg.max_level - depth, pg.max_level
        pset = front_end(pg, depth)
        string = fas_v_cycle_string(depth, pre, post, 14, cgc_relax_index=cgc)
        expression, _ = eval(string, pset.context, {})
        text = pg.generate_cycle_function(expression)
        out[name] = {"min_level": lo, "max_level": hi, "pre": pre, "post": post, "cgc_relax_index": cgc,
                     "template_smoother": True, "grammar": string, "layer4": text}
        print(name, len(text.splitlines()), "lines")
    inline = ProgramGeneratorFAS("FAS_2D_Basic", "Solution", "RHS", "Residual", "Approximation", "RestrictionNode",
                                 "CorrectionNode", "Laplace", "gamSten", "gen_mgCycle", "CGS", None)
    for name, depth, string in INLINE_CASES:
        if string is None:
            string = fas_w_like_string()
        pset = front_end(inline, depth)
        expression, _ = eval(string, pset.context, {})
        text = inline.generate_cycle_function(expression)
        out[name] = {"min_level": inline.max_level - depth, "max_level": inline.max_level, "template_smoother": False,
                     "grammar": string, "layer4": text}
        print(name, len(text.splitlines()), "lines")
    path = os.path.join(REPO, "tests", "golden", "fas_golden.json")
    with open(path, "w") as fh:
        json.dump(out, fh, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
