"""equations / operators / fields records for `generate_primitive_set`.

The reference fills these by parsing ExaStencils' debug L3 output
(/root/reference/evostencils/code_generation/parser.py:25-110).  Here they are built from
the problem specification.  When the reference package is importable the reference's own
record classes are used (so the unmodified `Optimizer` / `generate_primitive_set` accept
them, grammar/multigrid.py:15-71); otherwise light records with the same attributes.
"""
from __future__ import annotations

from . import stencils as st


class _OperatorInfo:
    def __init__(self, name, level, stencil, operator_type):
        self.name, self.level, self.stencil, self.operator_type = name, level, stencil, operator_type
        self._associated_field = None


class _EquationInfo:
    def __init__(self, name, level, expr_str, rhs_name, field):
        self.name, self.level, self.expr_str, self.rhs_name = name, level, expr_str, rhs_name
        self._associated_field = field

    @property
    def associated_field(self):
        return self._associated_field


def build_front_end(problem, min_level, max_level, parameters=None):
    params = dict(problem.parameters)
    params.update(parameters or {})
    try:
        import sympy
        from evostencils.grammar import multigrid as mg
        from evostencils.ir import base
        from evostencils.stencils import constant
        reference = True
    except Exception:
        reference = False
    dim = problem.dimension
    r_names = problem.restriction_names or [f"gen_restrictionForRes_{f}" for f in problem.field_names]
    p_names = problem.prolongation_names or [f"gen_prolongationForSol_{f}" for f in problem.field_names]
    if reference:
        fields = [sympy.Symbol(n) for n in problem.field_names]
        conv = lambda s: constant.Stencil(list(s.entries), s.dimension)
        kinds = (base.Operator, base.Restriction, base.Prolongation)
    else:
        fields = list(problem.field_names)
        conv = lambda s: s
        kinds = ("Operator", "Restriction", "Prolongation")
    equations, operators = [], []
    for level in range(min_level, max_level + 1):
        terms = problem.entry_terms(problem.spacing(level), params)
        seen = set()
        for i, row in enumerate(terms):
            lhs = []
            for j, entry in enumerate(row):
                for factor, name, sten in entry:
                    f = "" if factor is None else f"{factor!r} * "
                    lhs.append(f"( {f}{name}@{level} * {problem.field_names[j]}@{level} )")
                    if name not in seen:
                        seen.add(name)
                        info = (mg.OperatorInfo if reference else _OperatorInfo)(name, level, conv(sten), kinds[0])
                        operators.append(info)
            text = "( " + " + ".join(lhs) + f" ) == {problem.rhs_names[i]}@{level}"
            if reference:
                eq = mg.EquationInfo(problem.equation_names[i], level, text)
                eq._associated_field = fields[i]
            else:
                eq = _EquationInfo(problem.equation_names[i], level, text, problem.rhs_names[i], fields[i])
            equations.append(eq)
        for rn, pn in zip(r_names, p_names):
            cls = mg.OperatorInfo if reference else _OperatorInfo
            operators.append(cls(rn, level, conv(st.full_weighting(dim)), kinds[1]))
            operators.append(cls(pn, level, conv(st.multilinear_interpolation(dim)), kinds[2]))
    return equations, operators, fields
