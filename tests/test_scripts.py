"""scripts/ -- the B200 counterparts of the reference's entry-point scripts."""
import importlib.util
import os

import numpy as np
import pytest

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REPO, "scripts", name + ".py"))
    module = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(module)
    return module


def individual_file(tmp_path, levels, pre, post, omega_index, partitioning):
    from evostencils_b200.grammar import v_cycle_string
    path = tmp_path / "individual_0.txt"
    path.write_text(v_cycle_string(levels, pre, post, omega_index, partitioning) + "\n")
    return str(path)


def test_evolved_solver_script_emits_exaslang_and_needs_a_gpu(tmp_path):
    """Without a CUDA device the script still writes the ExaSlang hand-off, then fails loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    from evostencils_b200.engine import EngineError
    script = load("evaluate_evolved_solver")
    exa = tmp_path / "cycle.exa4"
    with pytest.raises(EngineError, match="no CPU fallback"):
        script.main([individual_file(tmp_path, 4, 2, 1, 21, "red_black"), "--exa", str(exa)])
    text = exa.read_text()
    assert "Function gen_mgCycle@9" in text and "solve locally" in text


def test_reference_solver_script_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    from evostencils_b200.engine import EngineError
    with pytest.raises(EngineError, match="no CPU fallback"):
        load("evaluate_reference_solver").main(["--nruns", "1"])


def test_fas_script_needs_a_gpu(tmp_path):
    import json
    import torch
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    from evostencils_b200.engine import EngineError
    script = load("evaluate_fas_solver")
    with pytest.raises(EngineError, match="no CPU fallback"):
        script.main(["--min-level", "3", "--max-level", "5", "--nruns", "1"])
    golden = json.load(open(os.path.join(REPO, "tests", "golden", "fas_golden.json")))["fas_inline_v_two_coarsenings"]
    path = tmp_path / "individual_0.txt"
    path.write_text(golden["grammar"])
    with pytest.raises(EngineError, match="no CPU fallback"):
        script.main(["--individual", str(path), "--min-level", "4", "--max-level", "6", "--nruns", "1"])


@pytest.mark.gpu
def test_fas_script_evaluates_an_individual(tmp_path, capsys):
    import json
    golden = json.load(open(os.path.join(REPO, "tests", "golden", "fas_golden.json")))["fas_inline_v_two_coarsenings"]
    path = tmp_path / "individual_0.txt"
    path.write_text(golden["grammar"])
    n, t = load("evaluate_fas_solver").main(["--individual", str(path), "--min-level", "4", "--max-level", "6", "--nruns", "2"])
    assert 3 < n <= 100 and t > 0 and "total no of iterations" in capsys.readouterr().out
    n2, t2 = load("evaluate_fas_solver").main(["--min-level", "3", "--max-level", "6", "--nruns", "2"])
    assert n2 == 13                                                   # oracle: FAS V(2,2) on levels 3..6


@pytest.mark.gpu
def test_reference_solver_script_matches_the_oracle(capsys):
    """Config 1 in small: the `.exa3` default solver (V(2,1) RB-GS, CG below) on levels 3..6."""
    from evostencils_b200 import lowering, problems
    from oracle import driver
    t, n = load("evaluate_reference_solver").main(["--min-level", "3", "--max-level", "6", "--nruns", "2"])
    out = capsys.readouterr().out
    assert "Average solving time:" in out and "Average number of iterations:" in out
    problem = problems.POISSON_2D
    operators = {level: lowering.operator_table(problem.system_operator(level, 0)) for level in range(3, 7)}
    program = lowering.Program(lowering.default_cycle_ops(problem, 6, 3), 3, 6, 1, 2, operators)
    _, hist, rho_ref, n_ref = driver.solve(problem, program)
    assert n == n_ref and t > 0


@pytest.mark.gpu
def test_evolved_solver_script_matches_the_oracle(tmp_path, capsys):
    from evostencils_b200 import lowering, problems
    from evostencils_b200.grammar import Grammar
    from oracle import driver
    path = individual_file(tmp_path, 3, 2, 2, 14, "single")
    t, rho, n = load("evaluate_evolved_solver").main([path, "--min-level", "3", "--max-level", "6",
                                                      "--evaluation-samples", "2"])
    assert "Convergence factor:" in capsys.readouterr().out
    problem = problems.POISSON_2D
    expression, _ = Grammar(problem, 3, 6, maximum_local_system_size=8).compile(open(path).read().strip())
    _, hist, rho_ref, n_ref = driver.solve(problem, lowering.lower(expression, 3, 6))
    assert n == n_ref
    assert abs(rho - rho_ref) <= 1e-9 * rho_ref
