"""Nonlinear full-approximation-scheme path: the reference's FAS example
`example_problems/FAS_2D_Basic` (-Laplace(u) + gam u e^u = f) on the GPU engine.

The reference emits Layer-4 ExaSlang for such a solver (`ProgramGeneratorFAS.traverse_graph`,
/root/reference/evostencils/code_generation/exastencils_FAS.py:99-319), pastes it into
`FAS_2D_Basic_template.exa4`, runs the ExaStencils compiler and the generated binary and parses
"total no of iterations" / "time to solution" (:369-426).  Here the statement sequence of the cycle
function -- `FAS_2D_Basic.exa4:210-269` is one such sequence, a V(2,2) cycle over three levels -- is
a list of `FasOp` records executed by the stage kernels of `csrc/mg_fas.cuh`; `solve` is the `Solve`
function of the `.exa4` (:118-187).  Not built yet: the grammar productions that make such cycles
(grammar/multigrid.py:275-299, 360-375) and the red-black / Picard smoother variants.
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass
from typing import List, Tuple

import numpy as np

from . import lowering
from .engine import Workspace

GAMMA = 10.0                  # Globals { Val gam : Real = 10.0 }  (.exa4:32-34)
SMOOTHER_WEIGHT = 0.8         # Smoother@all                        (.exa4:65-75)
CGS_SWEEPS = 30               # CGS@coarsest: repeat 30 times       (.exa4:58-63)
TARGET_REDUCTION = 1e-10      # repeat until res < 1e-10 * resStart (.exa4:147)
MAX_ITERATIONS = 100


@dataclass
class FasOp:
    kind: str                 # smooth | residual | restrict | subtract | prolong_add
    level: int                # level of the field that is written
    omega: float = 1.0
    sweeps: int = 1


def laplace_table(level: int) -> list:
    """`Stencil Laplace` of the `.exa4` (:18-24) as the engine's canonical operator table
    (offsets in memory order S, W, C, E, N)."""
    h2 = float(4 ** level)
    return [[[((0, -1), -h2), ((-1, 0), -h2), ((0, 0), 2.0 * h2 + 2.0 * h2), ((1, 0), -h2), ((0, 1), -h2)]]]


def right_hand_side(level: int, gamma: float = GAMMA) -> np.ndarray:
    """rhsFct of the enhanced test problem (.exa4:46-53); [y, x]."""
    n = 2 ** level + 1
    x = np.linspace(0.0, 1.0, n)[None, :]
    y = np.linspace(0.0, 1.0, n)[:, None]
    sol = (x ** 2 - x ** 3) * np.sin(3.0 * np.pi * y)
    return ((9.0 * np.pi ** 2 + gamma * np.exp(sol)) * (x ** 2 - x ** 3) + 6.0 * x - 2.0) * np.sin(3.0 * np.pi * y)


def exact_solution(level: int) -> np.ndarray:
    n = 2 ** level + 1
    x = np.linspace(0.0, 1.0, n)[None, :]
    y = np.linspace(0.0, 1.0, n)[:, None]
    return (x ** 2 - x ** 3) * np.sin(3.0 * np.pi * y)


def v_cycle_ops(max_level: int, min_level: int, pre: int = 2, post: int = 2, omega: float = SMOOTHER_WEIGHT,
                cgs_sweeps: int = CGS_SWEEPS, cgc_omega: float = 1.0) -> List[FasOp]:
    """The statement sequence of `gen_mgCycle@finest` for a FAS V(pre, post) cycle, in the order
    `traverse_graph` emits it (.exa4:210-269 for max_level 5, min_level 3, pre = post = 2; pinned on
    listings emitted by the unmodified reference for FAS grammar strings, tests/golden/fas_golden.json).
    `cgc_omega` = relaxation factor of the coarse-grid correction (`Solution += w * (P * Solution_c)`,
    exastencils_FAS.py:104-118; the hand-written sample has none)."""
    ops: List[FasOp] = []

    def visit(level):
        if level == min_level:
            ops.append(FasOp("smooth", level, omega, cgs_sweeps))            # CGS@coarsest
            return
        ops.append(FasOp("smooth", level, omega, pre))
        ops.append(FasOp("residual", level))
        ops.append(FasOp("restrict", level - 1))                             # approximation, solution and rhs of level-1
        visit(level - 1)
        ops.append(FasOp("subtract", level - 1))
        ops.append(FasOp("prolong_add", level, cgc_omega))
        ops.append(FasOp("smooth", level, omega, post))

    visit(max_level)
    return ops


class FasSolver:
    """FAS_2D_Basic on one GPU: level hierarchy, right-hand side of the enhanced test problem, zero
    initial guess; `solve` runs a cycle until the `.exa4`'s stopping rule holds."""

    def __init__(self, min_level: int, max_level: int, gamma: float = GAMMA, device: int = 0):
        self.min_level, self.max_level, self.gamma = min_level, max_level, gamma
        self.ws = Workspace(2, min_level, max_level, 1, "f64", device)
        for level in range(min_level, max_level + 1):
            self.ws.set_operator(level, laplace_table(level))
        self.reset()

    def reset(self):
        n = 2 ** self.max_level + 1
        self.ws.upload("rhs", self.max_level, 0, right_hand_side(self.max_level, self.gamma))
        self.ws.upload("sol", self.max_level, 0, np.zeros((n, n)))

    def run_cycle(self, ops: List[FasOp]):
        ws, gamma = self.ws, self.gamma
        for op in ops:
            if op.kind == "smooth":
                if op.sweeps > 0:
                    ws.fas_smooth(op.level, op.omega, gamma, op.sweeps)
            elif op.kind == "residual":
                ws.fas_residual(op.level, gamma)
            elif op.kind == "restrict":
                ws.fas_restrict(op.level, gamma)
            elif op.kind == "subtract":
                ws.fas_subtract_approximation(op.level)
            elif op.kind == "prolong_add":
                ws.exec_ops([lowering.Op("prolong_add", op.level, omega=op.omega, src="sol", dst="sol")])
            else:
                raise ValueError(f"unknown FAS op '{op.kind}'")

    def solve(self, ops: List[FasOp], target: float = TARGET_REDUCTION,
              max_iterations: int = MAX_ITERATIONS) -> Tuple[List[float], float]:
        """-> (residual history, solve time in ms): `Solve` of the `.exa4` (:118-187)."""
        history = [self.ws.fas_residual_norm(self.max_level, self.gamma)]
        t0 = time.perf_counter()
        while not (history[-1] < target * history[0] or len(history) - 1 >= max_iterations):
            self.run_cycle(ops)
            history.append(self.ws.fas_residual_norm(self.max_level, self.gamma))
            if not math.isfinite(history[-1]):
                break
        return history, (time.perf_counter() - t0) * 1e3

    def solution(self) -> np.ndarray:
        return self.ws.download("sol", self.max_level, 0)

    def close(self):
        self.ws.close()
