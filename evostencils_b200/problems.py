"""Problem definitions for the configurations of BASELINE.json.

The reference obtains these by running the ExaStencils compiler on the `.exa2/.exa3`
files and parsing its *debug* L3 output (code_generation/exastencils.py:93-103,
code_generation/parser.py:25-110).  Without java that is impossible, so the same
information is stated programmatically here, transcribed from

  * example_problems/Poisson/2D_FD_Poisson_fromL2.{exa2,exa3,knowledge}
  * example_problems/Poisson/3D_FD_Poisson_fromL2.{exa2,exa3,knowledge}
  * example_problems/LinearElasticity/2D_FD_LinearElasticity_fromL2.{exa2,knowledge}
  * example_problems/Helmholtz/2D_FD_Helmholtz_fromL3.{exa3,exa4,knowledge}

Grid convention (exastencils.py:97-103): level L has 2**L cells and 2**L+1 nodes per
dimension on the unit square/cube, h = 2**-L.  Arrays are indexed [i_{d-1}]...[i0] with
i0 (x) fastest.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np

from . import ir
from . import stencils as st


@dataclass
class SolverDefaults:
    """`generate solver ... with {...}` block of the `.exa3` (e.g. Poisson/2D...exa3:2-15)."""
    target_reduction: float = 1e-12
    max_iterations: int = 100
    num_pre: int = 2
    num_post: int = 1
    damping: float = 1.15
    coloring: str = "red_black"
    cgs: str = "CG"
    cgs_max_iterations: int = 1000
    cgs_target_reduction: float = 1e-12


@dataclass
class ProblemSpec:
    name: str
    dimension: int
    field_names: List[str]
    rhs_names: List[str]
    equation_names: List[str]
    min_level: int
    max_level: int
    dtype: str                                   # "f64" | "c128"
    # entry_terms(level_spacing, params) -> F x F matrix of lists of (factor, op_name, Stencil)
    entry_terms: Callable[[Sequence[float], Dict], List[List[list]]]
    rhs_function: Callable[..., List[np.ndarray]]          # (coords..., params) -> per-field arrays
    boundary_function: Callable[..., List[np.ndarray]]     # (coords..., params) -> per-field arrays
    solver: SolverDefaults = field(default_factory=SolverDefaults)
    outer: str = "mg"                            # "mg" | "bicgstab"
    boundary: str = "dirichlet"                  # "dirichlet" | "helmholtz_robin"
    parameters: Dict = field(default_factory=dict)
    restriction_names: Optional[List[str]] = None
    prolongation_names: Optional[List[str]] = None
    solver_iteration_limit: Optional[int] = 500  # scripts/optimize.py:32
    # {"kind": "exp_kappa", "kappa": <parameter name or number>}: the operator is -div(c grad u) with the
    # coefficient of stencils/gallery.py:87-185 evaluated at every node; `entry_terms` then gives the
    # stencil at the centre of the domain (structure + a representative for the IR, as the reference's
    # generate_stencil does for one position)
    variable_coefficient: Optional[Dict] = None

    @property
    def n_fields(self):
        return len(self.field_names)

    def variable_coefficient_spec(self, params=None):
        """-> {"kind", "kappa": float} with the parameter value substituted, or None."""
        if not self.variable_coefficient:
            return None
        p = dict(self.parameters)
        p.update(params or {})
        kappa = self.variable_coefficient["kappa"]
        return {"kind": self.variable_coefficient["kind"], "kappa": float(p[kappa] if isinstance(kappa, str) else kappa)}

    # -- numeric helpers ---------------------------------------------------------------
    def spacing(self, level):
        return tuple([1 / (2 ** level)] * self.dimension)

    def grids(self, level) -> List[ir.Grid]:
        return [ir.Grid(tuple([2 ** level] * self.dimension), self.spacing(level), level)
                for _ in self.field_names]

    def operator_stencils(self, level, params=None) -> List[List[st.Stencil]]:
        p = dict(self.parameters)
        p.update(params or {})
        terms = self.entry_terms(self.spacing(level), p)
        out = []
        for row in terms:
            out.append([])
            for entry in row:
                acc = None
                for factor, _name, sten in entry:
                    s = sten if factor is None else sten.scaled(factor)
                    acc = s if acc is None else acc + s
                out[-1].append(acc if acc is not None else st.null_stencil(self.dimension))
        return out

    def system_operator(self, level, depth, params=None) -> ir.SystemOperator:
        """Block operator `A_{depth}` whose entries are expression trees with the same
        structure `generate_operator_entries_from_equation` builds
        (grammar/multigrid.py:74-123): Addition(Scaling(f, Operator), ...)."""
        p = dict(self.parameters)
        p.update(params or {})
        grids = self.grids(level)
        terms = self.entry_terms(self.spacing(level), p)
        entries = []
        for i, row in enumerate(terms):
            entries.append([])
            for j, entry in enumerate(row):
                node = None
                for factor, name, sten in entry:
                    op = ir.Operator(name, grids[j], ir.ConstantStencilGenerator(sten))
                    term = op if factor is None else ir.Scaling(factor, op)
                    node = term if node is None else ir.Addition(node, term)
                entries[-1].append(node if node is not None else ir.ZeroOperator(grids[j]))
        return ir.SystemOperator(f"A_{depth}", entries)

    def node_coordinates(self, level):
        n = 2 ** level + 1
        axes = [np.arange(n, dtype=np.float64) * (1 / (2 ** level)) for _ in range(self.dimension)]
        # arrays are [i_{d-1}]...[i0]; return coordinate arrays (x, y[, z]) broadcastable to that
        mesh = np.meshgrid(*axes[::-1], indexing="ij")
        return tuple(mesh[::-1])  # x varies along the last axis

    def initial_fields(self, level, params=None):
        """(u0, f) per field on `level`: interior of u0 = 0, boundary = BC function
        (finest level only; coarser levels are error equations with zero boundary),
        f = RHS function (O2, O10, O11 of SURVEY.md 8c)."""
        p = dict(self.parameters)
        p.update(params or {})
        coords = self.node_coordinates(level)
        h = 1 / (2 ** level)
        np_dtype = np.complex128 if self.dtype == "c128" else np.float64
        bc = self.boundary_function(*coords, h=h, **p)
        rhs = self.rhs_function(*coords, h=h, **p)
        us, fs = [], []
        interior = tuple([slice(1, -1)] * self.dimension)
        for b, r in zip(bc, rhs):
            u = np.array(np.broadcast_to(b, coords[0].shape), dtype=np_dtype)
            u[interior] = 0
            f = np.array(np.broadcast_to(r, coords[0].shape), dtype=np_dtype)
            # boundary values of f are never read by any stage; zero them for determinism
            mask = np.ones_like(f, dtype=bool)
            mask[interior] = False
            f[mask] = 0
            us.append(u)
            fs.append(f)
        return us, fs


# ----------------------------------------------------------------------------------------
# Poisson 2D / 3D
# ----------------------------------------------------------------------------------------

def _poisson_terms(spacing, params):
    return [[[(None, "Laplace", st.poisson(spacing))]]]


def _poisson2d_rhs(x, y, h=None, **_):
    # RHS_u = PI**2 * cos(PI*x) - 4*PI**2 * sin(2*PI*y)   (Poisson/2D...exa2:7)
    return [np.pi ** 2 * np.cos(np.pi * x) - 4.0 * np.pi ** 2 * np.sin(2.0 * np.pi * y)]


def _poisson2d_bc(x, y, h=None, **_):
    # u on boundary = cos(PI*x) - sin(2*PI*y)              (Poisson/2D...exa2:5)
    return [np.cos(np.pi * x) - np.sin(2.0 * np.pi * y)]


def _poisson3d_rhs(x, y, z, h=None, **_):
    return [np.zeros(np.broadcast(x, y, z).shape)]          # Poisson/3D...exa2:9


def _poisson3d_bc(x, y, z, h=None, **_):
    # u@finest on boundary = x*x - 0.5*y*y - 0.5*z*z         (Poisson/3D...exa2:6)
    return [x * x - 0.5 * y * y - 0.5 * z * z]


POISSON_2D = ProblemSpec(
    name="2D_FD_Poisson_fromL2", dimension=2, field_names=["u"], rhs_names=["RHS_u"],
    equation_names=["solEq"], min_level=5, max_level=9, dtype="f64",
    entry_terms=_poisson_terms, rhs_function=_poisson2d_rhs, boundary_function=_poisson2d_bc,
    solver=SolverDefaults(damping=1.15))

POISSON_3D = ProblemSpec(
    name="3D_FD_Poisson_fromL2", dimension=3, field_names=["u"], rhs_names=["RHS_u"],
    equation_names=["solEq"], min_level=2, max_level=6, dtype="f64",
    entry_terms=_poisson_terms, rhs_function=_poisson3d_rhs, boundary_function=_poisson3d_bc,
    solver=SolverDefaults(damping=1.25))


# ----------------------------------------------------------------------------------------
# Variable-coefficient Poisson 2D / 3D: -div(c grad u) = f, c = exp(kappa * prod_d (x_d - x_d^2))
# (stencils/gallery.py:87-185).  The reference ships the stencil generators but no example problem for them;
# right-hand side and boundary values are those of the manufactured solution u = prod_d sin(pi x_d)
# (zero on the boundary), so that the discretisation error of the oracle can be checked against it.
# ----------------------------------------------------------------------------------------

def _varcoef_terms(spacing, params):
    centre = (0.5,) * len(spacing)
    return [[[(None, "A", st.variable_poisson(spacing, centre, params["kappa"]))]]]


def varcoef_solution(*coords):
    u = 1.0
    for x in coords:
        u = u * np.sin(np.pi * x)
    return u


def _varcoef_rhs(*coords, h=None, kappa=10.0, **_):
    # f = -div(c grad u) = c * (d pi^2 u - kappa * sum_a (prod_{b != a} (x_b - x_b^2)) (1 - 2 x_a) du/dx_a)
    d = len(coords)
    prod = 1.0
    for x in coords:
        prod = prod * (x - x * x)
    c = np.exp(kappa * prod)
    u = varcoef_solution(*coords)
    drift = 0.0
    for a in range(d):
        others = 1.0
        du = np.pi * np.cos(np.pi * coords[a])
        for b in range(d):
            if b != a:
                others = others * (coords[b] - coords[b] * coords[b])
                du = du * np.sin(np.pi * coords[b])
        drift = drift + others * (1.0 - 2.0 * coords[a]) * du
    return [c * (d * np.pi ** 2 * u - kappa * drift)]


def _varcoef_bc(*coords, h=None, **_):
    return [np.zeros(np.broadcast(*coords).shape)]


POISSON_2D_VARIABLE = ProblemSpec(
    name="2D_FD_Poisson_VariableCoefficients", dimension=2, field_names=["u"], rhs_names=["RHS_u"],
    equation_names=["solEq"], min_level=2, max_level=9, dtype="f64",
    entry_terms=_varcoef_terms, rhs_function=_varcoef_rhs, boundary_function=_varcoef_bc,
    solver=SolverDefaults(damping=1.15), parameters={"kappa": 10.0},
    variable_coefficient={"kind": "exp_kappa", "kappa": "kappa"})

POISSON_3D_VARIABLE = ProblemSpec(
    name="3D_FD_Poisson_VariableCoefficients", dimension=3, field_names=["u"], rhs_names=["RHS_u"],
    equation_names=["solEq"], min_level=2, max_level=6, dtype="f64",
    entry_terms=_varcoef_terms, rhs_function=_varcoef_rhs, boundary_function=_varcoef_bc,
    solver=SolverDefaults(damping=1.25), parameters={"kappa": 10.0},
    variable_coefficient={"kind": "exp_kappa", "kappa": "kappa"})


# ----------------------------------------------------------------------------------------
# Nonlinear FAS example: -Laplace(u) + gam u e^u = f   (example_problems/FAS_2D_Basic)
# ----------------------------------------------------------------------------------------

def _fas_solution(x, y):
    return (x ** 2 - x ** 3) * np.sin(3.0 * np.pi * y)                  # solFct, FAS_2D_Basic.exa4:54-56


def _fas_rhs(x, y, h=None, gam=10.0, **_):
    # rhsFct of the enhanced test problem (.exa4:46-53)
    return [((9.0 * np.pi ** 2 + gam * np.exp(_fas_solution(x, y))) * (x ** 2 - x ** 3) + 6.0 * x - 2.0)
            * np.sin(3.0 * np.pi * y)]


def _fas_bc(x, y, h=None, **_):
    return [np.zeros(np.broadcast(x, y).shape)]                         # Field Solution< global, BasicComm, 0.0 >


# the linear part only: the term gam e^u u is applied by the FAS stage kernels (csrc/mg_fas.cuh); names as
# ProgramGeneratorFAS registers them (exastencils_FAS.py:77-85); levels of the .knowledge file
FAS_2D_BASIC = ProblemSpec(
    name="FAS_2D_Basic", dimension=2, field_names=["u"], rhs_names=["RHS_u"],
    equation_names=["solEq"], min_level=6, max_level=10, dtype="f64",
    entry_terms=_poisson_terms, rhs_function=_fas_rhs, boundary_function=_fas_bc,
    solver=SolverDefaults(target_reduction=1e-10, max_iterations=100, damping=0.8, coloring="single"),
    parameters={"gam": 10.0}, restriction_names=["RestrictionNode"], prolongation_names=["ProlongationNode"])


# ----------------------------------------------------------------------------------------
# Linear elasticity 2D (two coupled fields)
# ----------------------------------------------------------------------------------------

def _elasticity_terms(spacing, params):
    lam, mu = params["lambda"], params["mu"]
    dxx = st.second_derivative(spacing, 0)
    dyy = st.second_derivative(spacing, 1)
    dxy = st.mixed_derivative_2d(spacing)
    lap = st.laplace(spacing)
    # uEq: (lambda+mu)*(dxx*u + dxy*v) + lambda*Laplace*u == RHS_u   (.exa2:45-47)
    # vEq: (lambda+mu)*(dxy*u + dyy*v) + lambda*Laplace*v == RHS_v   (.exa2:48-50)
    return [
        [[(lam + mu, "dxx", dxx), (lam, "Laplace", lap)], [(lam + mu, "dxy", dxy)]],
        [[(lam + mu, "dxy", dxy)], [(lam + mu, "dyy", dyy), (lam, "Laplace", lap)]],
    ]


def _elasticity_rhs(x, y, h=None, **_):
    z = np.zeros(np.broadcast(x, y).shape)
    return [z, z.copy()]


def _elasticity_bc(x, y, h=None, **_):
    # v on boundary = 4e-1 * sin(PI*x) * (1-x) * x * y   (.exa2:7)
    return [np.zeros(np.broadcast(x, y).shape), 4e-1 * np.sin(np.pi * x) * (1.0 - x) * x * y]


LINEAR_ELASTICITY_2D = ProblemSpec(
    name="2D_FD_LinearElasticity_fromL2", dimension=2, field_names=["u", "v"],
    rhs_names=["RHS_u", "RHS_v"], equation_names=["uEq", "vEq"], min_level=4, max_level=8,
    dtype="f64", entry_terms=_elasticity_terms, rhs_function=_elasticity_rhs,
    boundary_function=_elasticity_bc, solver=SolverDefaults(damping=1.25),
    parameters={"lambda": 195.0, "mu": 130.0})


# ----------------------------------------------------------------------------------------
# Helmholtz 2D (complex shifted-Laplacian preconditioner inside BiCGStab)
# ----------------------------------------------------------------------------------------

def _helmholtz_terms(spacing, params):
    # the multigrid cycle acts on `M * u == f` (PrecEq, Helmholtz...exa3:71-73)
    return [[[(None, "M", st.helmholtz_2d(spacing, params["k"], params["shift"]))]]]


def _helmholtz_rhs(x, y, h=None, **_):
    # B-spline approximation of the Dirac pulse (Helmholtz...exa3:24)
    a = np.maximum(0.0, -(np.abs(x - 0.5) - h) / h ** 2)
    b = np.maximum(0.0, -(np.abs(y - 0.5) - h) / h ** 2)
    return [(a * b).astype(np.complex128)]


def _helmholtz_bc(x, y, h=None, **_):
    return [np.zeros(np.broadcast(x, y).shape, dtype=np.complex128)]


HELMHOLTZ_2D = ProblemSpec(
    name="2D_FD_Helmholtz_fromL3", dimension=2, field_names=["u"], rhs_names=["f"],
    equation_names=["PrecEq"], min_level=3, max_level=7, dtype="c128",
    entry_terms=_helmholtz_terms, rhs_function=_helmholtz_rhs, boundary_function=_helmholtz_bc,
    solver=SolverDefaults(target_reduction=1e-7, max_iterations=10000, damping=0.6,
                          cgs="BiCGStab", cgs_max_iterations=1000, cgs_target_reduction=1e-6),
    outer="bicgstab", boundary="helmholtz_robin",
    parameters={"k": 80.0, "shift": 1.0 + 0.5j},
    restriction_names=["Restriction_u"], prolongation_names=["Prolongation_u"],
    solver_iteration_limit=10000)


REGISTRY = {p.name: p for p in (POISSON_2D, POISSON_3D, LINEAR_ELASTICITY_2D, HELMHOLTZ_2D, FAS_2D_BASIC,
                                POISSON_2D_VARIABLE, POISSON_3D_VARIABLE)}


def get_problem(name: str) -> ProblemSpec:
    try:
        return REGISTRY[name]
    except KeyError:
        raise KeyError(f"unknown problem '{name}'; known: {sorted(REGISTRY)}") from None
