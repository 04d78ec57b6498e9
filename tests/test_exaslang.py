"""The emitted ExaSlang listing is byte-identical to the reference's `generate_cycle_function`
output (exastencils.py:318-336) for the 11 golden individuals (tests/golden/_make)."""
import json
import os

import pytest

from evostencils_b200 import problems
from evostencils_b200.exaslang import emit_cycle_function
from evostencils_b200.grammar import Grammar

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "lowering_golden.json")))


@pytest.mark.parametrize("case", sorted(GOLDEN))
def test_listing_is_byte_identical(case):
    g = GOLDEN[case]
    problem = problems.get_problem(g["problem"])
    grammar = Grammar(problem, g["min_level"], g["max_level"], maximum_local_system_size=g["maximum_local_system_size"])
    expression, _ = grammar.compile(g["grammar"])
    text = emit_cycle_function(problem, expression, g["min_level"], g["max_level"], g["max_level"])
    assert text == g["exaslang"]


def test_generator_method_matches():
    from evostencils_b200.generator import B200ProgramGenerator
    g = GOLDEN["poisson2d_v21_rbgs"]
    problem = problems.get_problem(g["problem"])
    grammar = Grammar(problem, g["min_level"], g["max_level"])
    expression, _ = grammar.compile(g["grammar"])
    gen = B200ProgramGenerator.__new__(B200ProgramGenerator)
    gen._problem, gen._cycle_name, gen._parameters = problem, "gen_mgCycle", {}
    text = gen.generate_cycle_function(expression, None, g["min_level"], g["max_level"], g["max_level"])
    assert text == g["exaslang"]
    expression2, _ = grammar.compile(g["grammar"])
    assert emit_cycle_function(problem, expression2, g["min_level"], g["max_level"], g["max_level"]) == text
