"""ORACLE (test infrastructure, never on the product path) -- PARITY UNPINNED.

numpy fp64 / complex128 execution of the op list produced by
`evostencils_b200.lowering.lower()`, i.e. a CPU restatement of what the
ExaStencils-generated solver does for the same cycle.  The arithmetic of the reference
lives in a third-party dependency that is not under /root/reference: ExaStencils v1.1
(README.md:21-32; call sites code_generation/exastencils.py:381-415, 417-443) needs
java + sbt, neither of which exists in this image, and the reference holds no golden
residual histories or known-answer vectors for this path.  The oracle
is therefore pinned only structurally: the op *sequence* is checked against text emitted
by the unmodified reference (tests/golden/lowering_golden.json), the operator values
against the coefficients that text prints, and the numerics against dense-matrix
identities, textbook convergence factors and the analytic solution
(tests/test_oracle.py).  "parity unpinned" at the ExaStencils boundary.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.

Semantics (decisions O1-O11 of SURVEY.md 8c, refined in DESIGN.md "Numerical contract"):
  * node-centred grids, n = 2**L + 1 per dimension, arrays [i_{d-1}]...[i0], x fastest;
  * all loops over interior nodes; Dirichlet boundary values are never written
    (Robin columns of the Helmholtz problem are slaved after every write of a solution field);
  * A u is accumulated in canonical stencil order (block column j ascending, offsets in
    memory order), plain multiply/add, no FMA contraction;
  * local solve: u_s += sum_t W[s][t] * r_t with W = omega * inv(L_tile), r = f - A u;
    Jacobi reads old values, red-black updates colour (sum i_d) % 2 == 0 first, in place;
  * restriction = full weighting (x) [1/4 1/2 1/4], prolongation = multilinear
    (x) [1/2 1 1/2], coarse node I <-> fine node 2I;
  * coarse-grid solve: CG (real) / BiCGStab as written in Helmholtz/...exa3:396-433
    (complex), zero initial guess, dot products in the engine's lane-strided butterfly order;
  * outer iteration, convergence factor and iteration count as parse_output reads them
    (exastencils.py:539-584).
"""
from __future__ import annotations

import itertools
import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np

DOT_LANES = 256          # thread count of the engine's coarse-solver CTA


# ----------------------------------------------------------------------------------------
# helpers
# ----------------------------------------------------------------------------------------

def _shift(a: np.ndarray, off: Sequence[int]) -> np.ndarray:
    """View of `a` at interior nodes displaced by `off` (off[0] = x = last axis)."""
    d = a.ndim
    sl = []
    for axis in range(d):
        o = off[d - 1 - axis]
        n = a.shape[axis]
        sl.append(slice(1 + o, n - 1 + o))
    return a[tuple(sl)]


def _interior(a: np.ndarray):
    return a[tuple([slice(1, -1)] * a.ndim)]


# ----------------------------------------------------------------------------------------
# complex arithmetic, spelled out
# ----------------------------------------------------------------------------------------
# The engine evaluates complex products / quotients / roots with the textbook formulas in separate
# IEEE operations (csrc/mg_common.cuh t_mul / t_div, compiled with --fmad=false).  numpy's and
# CPython's complex multiply may contract to FMA (AVX2/AVX-512 loops) and both divide by Smith's
# method, so `a * b` on complex128 is NOT bit-reproducible across machines.  Preconditioned BiCGStab
# at k = 80 amplifies last-bit differences into different iteration counts, and `north_star` asks
# for identical counts -- so the oracle states every complex operation in real arithmetic, in the
# operand order of the engine.  Real data takes the plain numpy path.

def _is_complex(x) -> bool:
    return isinstance(x, complex) or np.iscomplexobj(x)


def _pack(re, im):
    out = np.empty(np.shape(re), dtype=np.complex128)
    out.real = re
    out.imag = im
    return out if out.shape else complex(out)


def mul(a, b):
    """a * b.  Complex: (a.re b.re - a.im b.im, a.re b.im + a.im b.re), four products, one
    subtraction, one addition, no FMA (t_mul); real operands are promoted with a zero imaginary part
    like the engine's operator tables."""
    if not (_is_complex(a) or _is_complex(b)):
        return a * b
    ar, ai = np.real(a), np.imag(a)
    br, bi = np.real(b), np.imag(b)
    return _pack(ar * br - ai * bi, ar * bi + ai * br)


def scale(s, a):
    """real scalar * field (t_scale): component-wise for complex data."""
    if not _is_complex(a):
        return s * a
    return _pack(s * np.real(a), s * np.imag(a))


def div(a, b):
    """a / b for scalars.  Complex: ((a.re b.re + a.im b.im) / d, (a.im b.re - a.re b.im) / d) with
    d = b.re^2 + b.im^2 (t_div)."""
    if not (_is_complex(a) or _is_complex(b)):
        return a / b
    ar, ai, br, bi = float(np.real(a)), float(np.imag(a)), float(np.real(b)), float(np.imag(b))
    d = br * br + bi * bi
    with np.errstate(all="ignore"):
        return complex(np.float64(ar * br + ai * bi) / np.float64(d), np.float64(ai * br - ar * bi) / np.float64(d))


def modulus(z) -> float:
    """|z| = sqrt(re^2 + im^2) (t_abs)."""
    if not _is_complex(z):
        return abs(float(z))
    re, im = float(np.real(z)), float(np.imag(z))
    return math.sqrt(re * re + im * im)


def root(z):
    """Principal square root: r = |z|, (sqrt((r + re)/2), sign(im) sqrt((r - re)/2)) (t_sqrt)."""
    if not _is_complex(z):
        return math.sqrt(z) if z >= 0 else float("nan")
    re, im = float(np.real(z)), float(np.imag(z))
    r = modulus(z)
    if not math.isfinite(r):
        return complex(float("nan"), float("nan"))
    sr = math.sqrt(0.5 * (r + re))
    si = math.sqrt(max(0.0, 0.5 * (r - re)))
    return complex(sr, -si if im < 0 else si)


def butterfly_sum(v: np.ndarray) -> float:
    """Sum of `DOT_LANES` lane partials exactly as the coarse-solver CTA forms it:
    xor-butterfly (16, 8, 4, 2, 1) inside each warp, then the same over the warp sums."""
    v = np.asarray(v).reshape(DOT_LANES // 32, 32)
    w = v
    for off in (16, 8, 4, 2, 1):
        w = w[:, :off] + w[:, off:2 * off]
    w = w[:, 0]
    k = w.shape[0]
    while k > 1:
        k //= 2
        w = w[:k] + w[k:2 * k]
    return w[0]


def lane_dot(a: np.ndarray, b: np.ndarray):
    """dot(a, b) without conjugation over flat vectors, lane-strided partials + butterfly."""
    n = a.size
    pad = (-n) % DOT_LANES
    prod = mul(a.reshape(-1), b.reshape(-1))
    if pad:
        prod = np.concatenate([prod, np.zeros(pad, dtype=prod.dtype)])
    rows = prod.reshape(-1, DOT_LANES)
    part = rows[0].copy()
    for r in rows[1:]:
        part = part + r
    return butterfly_sum(part)


GRID_DOT_BLOCKS = 1184   # CTAs of the engine's grid-wide dot / norm kernels (8 x 148 SMs), 256 threads each


def grid_dot(a: np.ndarray, b: np.ndarray):
    """sum over the interior of a * b (no conjugation) in the order of the engine's two-stage
    reduction (csrc/mg_outer.cuh k_outer_dot + k_outer_dot_final; 2D fields): CTA c of
    min(n, 1184) owns the rows c, c + grid, ...; thread t of 256 accumulates x = 1 + t, 257 + t, ...
    of each of them in turn; butterfly over the CTA; then thread t adds the CTA sums t, t + 256, ...
    and the same butterfly gives the total."""
    n = a.shape[0]
    assert a.ndim == 2, "grid_dot mirrors the 2D kernel"
    blocks = min(n, GRID_DOT_BLOCKS)
    prod = mul(a, b)
    lanes = np.zeros((blocks, DOT_LANES), dtype=prod.dtype)
    for first in range(0, n, blocks):                     # one row per CTA and round
        rows = np.arange(first, min(first + blocks, n))
        rows = rows[(rows >= 1) & (rows <= n - 2)]
        if rows.size == 0:
            continue
        for x0 in range(1, n - 1, DOT_LANES):
            x1 = min(x0 + DOT_LANES, n - 1)
            lanes[rows - first, : x1 - x0] = lanes[rows - first, : x1 - x0] + prod[rows, x0:x1]
    partial = np.array([butterfly_sum(lanes[c]) for c in range(blocks)], dtype=prod.dtype)
    pad = (-blocks) % DOT_LANES
    if pad:
        partial = np.concatenate([partial, np.zeros(pad, dtype=prod.dtype)])
    rounds = partial.reshape(-1, DOT_LANES)
    acc = np.zeros(DOT_LANES, dtype=prod.dtype)
    for r in rounds:
        acc = acc + r
    return butterfly_sum(acc)


@dataclass
class BoundaryCondition:
    kind: str = "dirichlet"               # "dirichlet" | "helmholtz_robin"
    k: float = 0.0

    def factor(self, h):
        # u_0 = u_1 / (1 - i k h)   (Helmholtz/...exa4:25-59)
        return div(complex(1.0, 0.0), complex(1.0, -self.k * h))


class Hierarchy:
    """Field storage of one solver instance (CycleStorage, exastencils.py:14-26)."""

    def __init__(self, dimension, n_fields, min_level, max_level, dtype=np.float64,
                 bc: Optional[BoundaryCondition] = None):
        self.dimension = dimension
        self.n_fields = n_fields
        self.min_level = min_level
        self.max_level = max_level
        self.dtype = dtype
        self.bc = bc or BoundaryCondition()
        self.sol: Dict[int, List[np.ndarray]] = {}
        self.rhs: Dict[int, List[np.ndarray]] = {}
        self.res: Dict[int, List[np.ndarray]] = {}
        self.cor: Dict[int, List[np.ndarray]] = {}
        for level in range(min_level, max_level + 1):
            shape = (2 ** level + 1,) * dimension
            self.sol[level] = [np.zeros(shape, dtype) for _ in range(n_fields)]
            self.rhs[level] = [np.zeros(shape, dtype) for _ in range(n_fields)]
            self.res[level] = [np.zeros(shape, dtype) for _ in range(n_fields)]
            # correction field aliases the solution field below the finest level
            self.cor[level] = self.sol[level] if level < max_level else [np.zeros(shape, dtype) for _ in range(n_fields)]
        self.operators: Dict[int, list] = {}
        self.variable_levels = set()          # levels whose table holds value arrays (variable_poisson_table)
        self._colour_masks: Dict[int, List[np.ndarray]] = {}

    def field(self, kind, level):
        return {"sol": self.sol, "rhs": self.rhs, "res": self.res, "cor": self.cor}[kind][level]

    def colour_mask(self, level, colour):
        if level not in self._colour_masks:
            n = 2 ** level + 1
            idx = np.indices((n - 2,) * self.dimension).sum(axis=0) + self.dimension  # interior starts at 1
            self._colour_masks[level] = [(idx % 2) == 0, (idx % 2) == 1]
        return self._colour_masks[level][colour]

    def apply_bc(self, level, arrays):
        if self.bc.kind != "helmholtz_robin":
            return
        c = self.bc.factor(1.0 / 2 ** level)
        for u in arrays:
            u[:, 0] = mul(u[:, 1], c)
            u[:, -1] = mul(u[:, -2], c)
            u[0, :] = 0
            u[-1, :] = 0


# ----------------------------------------------------------------------------------------
# stages
# ----------------------------------------------------------------------------------------

def apply_operator(table, fields: List[np.ndarray], row: int):
    """(A u)_row at interior nodes; canonical accumulation order."""
    acc = None
    for j, entries in enumerate(table[row]):
        for off, val in entries:
            operand = _shift(fields[j], off)
            term = mul(complex(val), operand) if np.iscomplexobj(operand) else val * operand
            acc = term if acc is None else acc + term
    if acc is None:
        acc = np.zeros_like(_interior(fields[0]))
    return acc


def variable_poisson_table(dimension: int, level: int, kappa: float):
    """Operator table of -div(c grad u), c = exp(kappa * prod_d (x_d - x_d^2)), with one VALUE ARRAY
    (interior nodes) per stencil offset -- `Operator A from Stencil` of
    Poisson2DVariableCoefficients / Poisson3DVariableCoefficients.generate_exa3
    (stencils/gallery.py:113-131, 163-185), evaluated at every node as the generated solver does:
    getCoefficient at nodePosition +- 0.5 * gridWidth (:87-90, 137-140), off-centre value
    (-1.0 * c) / (h * h), centre ((cE + cW) / hh) + ((cN + cS) / hh) [+ ((cT + cB) / hh)].
    Entries in memory order of the offsets, like every other table."""
    n = 2 ** level + 1
    hw = 1.0 / 2 ** level
    hh = hw * hw
    pos = np.arange(1, n - 1, dtype=np.float64) * hw                     # interior node positions, exact
    shape = [1] * dimension

    def axis_array(values, x_axis):                                      # x_axis 0 = x = last array axis
        sh = list(shape)
        sh[dimension - 1 - x_axis] = values.size
        return values.reshape(sh)

    def coefficient(shift_axis, sign):
        prod = None
        for a in range(dimension):                                       # x, y[, z]: the reference's grouping
            p = pos + sign * (0.5 * hw) if a == shift_axis else pos
            t = axis_array(p - (p * p), a)
            prod = t if prod is None else prod * t
        full = np.broadcast_to(prod, (n - 2,) * dimension)
        return np.exp(kappa * full)

    faces = {}
    centre = None
    for a in range(dimension):
        cp, cm = coefficient(a, +1), coefficient(a, -1)
        faces[(a, +1)], faces[(a, -1)] = cp, cm
        term = (cp + cm) / hh
        centre = term if centre is None else centre + term
    entries = []
    for a in reversed(range(dimension)):                                 # memory order: slowest axis first
        off = [0] * dimension
        off[a] = -1
        entries.append((tuple(off), (-1.0 * faces[(a, -1)]) / hh))
    entries.append(((0,) * dimension, centre))
    for a in range(dimension):
        off = [0] * dimension
        off[a] = +1
        entries.append((tuple(off), (-1.0 * faces[(a, +1)]) / hh))
    return [[entries]]


def residual_interior(h: Hierarchy, level, row, sol=None):
    sol = h.sol[level] if sol is None else sol
    return _interior(h.rhs[level][row]) - apply_operator(h.operators[level], sol, row)


def op_residual(h: Hierarchy, level):
    for i in range(h.n_fields):
        _interior(h.res[level][i])[...] = residual_interior(h, level, i)


def op_zero(h: Hierarchy, level, kind="sol"):
    for a in h.field(kind, level):
        a[...] = 0


def op_restrict(h: Hierarchy, coarse_level, src_kind="res"):
    """rhs[coarse](I) = sum_o w_o src[fine](2I + o), offsets in memory order."""
    d = h.dimension
    w1 = {-1: 0.25, 0: 0.5, 1: 0.25}
    for i in range(h.n_fields):
        fine = h.field(src_kind, coarse_level + 1)[i]
        nc = 2 ** coarse_level + 1
        acc = None
        for off in itertools.product((-1, 0, 1), repeat=d):       # off[0] slowest axis
            w = 1.0
            for c in off:
                w *= w1[c]
            sl = tuple(slice(2 + c, 2 * (nc - 1) - 2 + c + 1, 2) for c in off)
            term = scale(w, fine[sl])
            acc = term if acc is None else acc + term
        _interior(h.rhs[coarse_level][i])[...] = acc


def prolongation_interior(coarse: np.ndarray, d: int):
    """(P e) at the interior fine nodes: 2**-k * (sum of the 2**k enclosing coarse nodes,
    memory order), k = number of odd fine coordinates."""
    nc = coarse.shape[0]
    nf = 2 * (nc - 1) + 1
    out = np.zeros((nf,) * d, dtype=coarse.dtype)
    for parity in itertools.product((0, 1), repeat=d):              # parity[0] = slowest axis
        k = sum(parity)
        dst = tuple(slice(p, nf, 2) for p in parity)
        acc = None
        for corner in itertools.product(*[((0, 1) if p else (0,)) for p in parity]):
            src = tuple(slice(c, nc - 1 + c) if p else slice(0, nc) for p, c in zip(parity, corner))
            term = coarse[src]
            acc = term if acc is None else acc + term
        out[dst] = scale(0.5 ** k, acc) if k else acc
    return _interior(out)


def op_prolong_add(h: Hierarchy, fine_level, omega, src_kind="sol"):
    for i in range(h.n_fields):
        coarse = h.field(src_kind, fine_level - 1)[i]
        _interior(h.sol[fine_level][i])[...] += scale(omega, prolongation_interior(coarse, h.dimension))
    h.apply_bc(fine_level, h.sol[fine_level])


def op_prolong(h: Hierarchy, fine_level, src_kind="sol"):
    for i in range(h.n_fields):
        coarse = h.field(src_kind, fine_level - 1)[i]
        _interior(h.cor[fine_level][i])[...] = prolongation_interior(coarse, h.dimension)


def inverse_block(matrix: np.ndarray, omega) -> np.ndarray:
    """omega * inv(L) -- closed form for 1x1 and 2x2 (shared, digit for digit, with the
    product's host code so that point smoothers are bit-identical), LAPACK otherwise."""
    m = matrix.shape[0]
    if m == 1:
        return np.array([[omega / matrix[0, 0]]], dtype=matrix.dtype)
    if m == 2:
        a, b, c, d = matrix[0, 0], matrix[0, 1], matrix[1, 0], matrix[1, 1]
        det = a * d - b * c
        return np.array([[omega * (d / det), omega * (-b / det)],
                         [omega * (-c / det), omega * (a / det)]], dtype=matrix.dtype)
    return omega * np.linalg.inv(matrix)


def op_local_solve(h: Hierarchy, stmt):
    level = stmt.level
    d = h.dimension
    n = 2 ** level + 1
    sol = h.sol[level]
    matrix = np.array(stmt.matrix, dtype=h.dtype)
    period = tuple(stmt.period)
    slots = list(stmt.slots)
    m = len(slots)
    if level in h.variable_levels:
        # variable-coefficient operator: the local 1x1 system has the node's own centre value
        assert m == 1 and all(p == 1 for p in period), "point smoothers only on variable-coefficient operators"
        diag = [v for off, v in h.operators[level][0][0] if not any(off)][0]
        W = [[stmt.omega / diag]]
    elif all(p == 1 for p in period):
        W = inverse_block(matrix, stmt.omega)
    if all(p == 1 for p in period):
        colours = (0, 1) if stmt.colored else (None,)
        for colour in colours:
            r = [residual_interior(h, level, f) for f, _ in slots]
            for s, (f, _) in enumerate(slots):
                delta = None
                for t in range(m):
                    term = mul(W[s][t], r[t])
                    delta = term if delta is None else delta + term
                target = _interior(sol[f])
                if colour is None:
                    target[...] = target + delta
                else:
                    mask = h.colour_mask(level, colour)
                    target[mask] = target[mask] + delta[mask]
            h.apply_bc(level, sol)
        return
    # tile smoother (Jacobi only): all residuals from the old iterate
    assert stmt.jacobi and not stmt.colored
    r_full = {}
    for f in {f for f, _ in slots}:
        a = np.zeros((n,) * d, dtype=h.dtype)
        _interior(a)[...] = residual_interior(h, level, f)
        r_full[f] = a
    # per-dimension tile classes: tiles with the same validity pattern of their positions
    classes = []
    for axis_x in range(d):                                         # axis_x = 0 is x
        b = period[axis_x]
        n_tiles = (n - 1) // b + 1
        by_mask: Dict[tuple, list] = {}
        for t in range(n_tiles):
            mask = tuple(1 <= t * b + p <= n - 2 for p in range(b))
            by_mask.setdefault(mask, []).append(t)
        classes.append(by_mask)
    for combo in itertools.product(*[list(c.items()) for c in classes]):
        masks = [c[0] for c in combo]                               # per x-dimension
        tiles = [np.array(c[1]) for c in combo]
        valid = [all(masks[ax][pos[ax]] for ax in range(d)) for _, pos in slots]
        if not any(valid):
            continue
        vidx = [s for s in range(m) if valid[s]]
        W = inverse_block(matrix[np.ix_(vidx, vidx)], stmt.omega)
        gathered = []
        index = []
        for s in vidx:
            f, pos = slots[s]
            ix = np.ix_(*[tiles[ax] * period[ax] + pos[ax] for ax in reversed(range(d))])
            index.append((f, ix))
            gathered.append(r_full[f][ix])
        for a, s in enumerate(vidx):
            delta = None
            for b_, t in enumerate(vidx):
                term = mul(W[a, b_], gathered[b_])
                delta = term if delta is None else delta + term
            f, ix = index[a]
            sol[f][ix] = sol[f][ix] + delta
    h.apply_bc(level, sol)


# ----------------------------------------------------------------------------------------
# coarse-grid solvers
# ----------------------------------------------------------------------------------------

def _flat_interior(arrays):
    return np.concatenate([_interior(a).reshape(-1) for a in arrays])


def _apply_flat(h: Hierarchy, level, arrays):
    return np.concatenate([apply_operator(h.operators[level], arrays, i).reshape(-1) for i in range(h.n_fields)])


def _scatter_interior(arrays, flat):
    k = 0
    for a in arrays:
        it = _interior(a)
        it[...] = flat[k:k + it.size].reshape(it.shape)
        k += it.size


def op_coarse_cg(h: Hierarchy, level, max_its=1000, reduction=1e-12, rhs=None, into=None):
    """Plain CG from a zero guess, stop at ||r|| <= reduction * ||r0|| (Poisson/2D...exa3:12-14).
    `rhs` / `into`: right-hand side and destination arrays other than the level's rhs / sol fields
    (the Krylov smoother solves the residual equation into a scratch field)."""
    x = h.sol[level] if into is None else into
    for a in x:
        a[...] = 0
    work = [np.zeros_like(a) for a in x]                             # p with zero Dirichlet rim
    r = _flat_interior(h.rhs[level] if rhs is None else rhs).copy()
    p = r.copy()
    xs = np.zeros_like(r)
    rs = lane_dot(r, r)
    r0 = modulus(root(rs))
    if r0 == 0.0:
        return 0
    its = 0
    for its in range(1, max_its + 1):
        _scatter_interior(work, p)
        ap = _apply_flat(h, level, work)
        alpha = div(rs, lane_dot(p, ap))
        xs = xs + mul(alpha, p)
        r = r - mul(alpha, ap)
        rs_new = lane_dot(r, r)
        if modulus(root(rs_new)) <= reduction * r0:
            break
        beta = div(rs_new, rs)
        p = r + mul(beta, p)
        rs = rs_new
    _scatter_interior(x, xs)
    return its


def op_coarse_bicgstab(h: Hierarchy, level, max_its=1000, reduction=1e-6, rhs=None, into=None):
    """Unpreconditioned BiCGStab exactly as Helmholtz/2D_FD_Helmholtz_fromL3.exa3:396-433
    (dot without conjugation; p and s carry the Robin boundary, :384-391).  `rhs` / `into` as op_coarse_cg."""
    x = h.sol[level] if into is None else into
    for a in x:
        a[...] = 0
    wp = [np.zeros_like(a) for a in x]
    ws = [np.zeros_like(a) for a in x]
    xs = np.zeros_like(_flat_interior(x))
    res = _flat_interior(h.rhs[level] if rhs is None else rhs) - _apply_flat(h, level, x)
    cur = root(lane_dot(res, res))
    init = modulus(cur)
    if cur == 0.0:
        return 0
    one = 1.0 + 0j if np.iscomplexobj(res) else 1.0
    alpha = beta = rho_new = omega = one
    res_hat = res.copy()
    nu = np.zeros_like(res)
    p = np.zeros_like(res)
    its = 0
    for its in range(1, max_its + 1):
        rho = rho_new
        rho_new = lane_dot(res_hat, res)
        beta = mul(div(rho_new, rho), div(alpha, omega))
        p = res + mul(beta, p - mul(omega, nu))
        _scatter_interior(wp, p)
        h.apply_bc(level, wp)
        nu = _apply_flat(h, level, wp)
        alpha = div(rho_new, lane_dot(res_hat, nu))
        hh = xs + mul(alpha, p)
        s = res - mul(alpha, nu)
        _scatter_interior(ws, s)
        h.apply_bc(level, ws)
        t = _apply_flat(h, level, ws)
        omega = div(lane_dot(t, s), lane_dot(t, t))
        xs = hh + mul(omega, s)
        res = s - mul(omega, t)
        cur = root(lane_dot(res, res))
        if modulus(cur) < reduction * init:
            break
    _scatter_interior(x, xs)
    h.apply_bc(level, x)
    return its


def op_krylov(h: Hierarchy, level, method, iterations, reduction):
    """`<Name>_<n>@level()` in the place of a smoother (code_generation/exastencils.py:758-768;
    ir/krylov_subspace.py:4-45): at most n iterations of the Krylov method on A x = rhs@level starting from
    the current solution.  The reference takes the function body from the ExaStencils-generated coarse-grid
    solver of that level (:1025-1086, marked "either fix or remove" there; its grammar never produces the
    node), so the statement is restated in the engine's terms: r = rhs - A x into the residual field, the
    coarse-grid solver's iteration on A e = r from e = 0 (n iterations at most, the `.exa3` block's
    reduction target as early exit), x += e on interior nodes.  Equivalent to starting the method from x."""
    op_residual(h, level)
    scratch = [np.zeros_like(a) for a in h.sol[level]]
    if method == "ConjugateGradient":
        op_coarse_cg(h, level, iterations, reduction, rhs=h.res[level], into=scratch)
    elif method == "BiCGStab":
        op_coarse_bicgstab(h, level, iterations, reduction, rhs=h.res[level], into=scratch)
    else:
        raise NotImplementedError(method)
    for u, e in zip(h.sol[level], scratch):
        _interior(u)[...] = _interior(u) + _interior(e)
    h.apply_bc(level, h.sol[level])


# ----------------------------------------------------------------------------------------
# op-list interpreter and outer solver
# ----------------------------------------------------------------------------------------

@dataclass
class SolverSettings:
    target_reduction: float = 1e-12
    max_iterations: int = 100
    cgs: str = "CG"
    cgs_max_iterations: int = 1000
    cgs_target_reduction: float = 1e-12
    default_cycle: Optional[callable] = None      # level -> op list, for `cgs` above min_level


def run_ops(h: Hierarchy, ops, settings: SolverSettings):
    for op in ops:
        kind = getattr(op, "kind")
        if kind == "smooth":
            op_local_solve(h, op)
        elif kind == "residual":
            op_residual(h, op.level)
        elif kind == "restrict":
            op_restrict(h, op.level, op.src or "res")
        elif kind == "residual_restrict":
            op_residual(h, op.level + 1)
            op_restrict(h, op.level, "res")
        elif kind == "zero":
            op_zero(h, op.level, op.dst or "sol")
        elif kind == "prolong_add":
            op_prolong_add(h, op.level, op.omega, op.src or "sol")
        elif kind == "prolong":
            op_prolong(h, op.level, op.src or "sol")
        elif kind == "krylov":
            op_krylov(h, op.level, op.method, op.iterations, settings.cgs_target_reduction)
        elif kind == "cgs":
            if op.level == h.min_level:
                if settings.cgs == "CG":
                    op_coarse_cg(h, op.level, settings.cgs_max_iterations, settings.cgs_target_reduction)
                else:
                    op_coarse_bicgstab(h, op.level, settings.cgs_max_iterations, settings.cgs_target_reduction)
            else:
                inner = op.inner if op.inner else settings.default_cycle(op.level)
                op_zero(h, op.level, "sol")
                run_ops(h, inner, settings)
        else:
            raise NotImplementedError(kind)


def residual_norm(h: Hierarchy, level=None):
    """sqrt(sum_interior r**2) over all fields, no h scaling (FAS_2D_Basic.exa4:83-89);
    complex: sqrt(dot(r, r)) without conjugation, then |.| (Helmholtz/...exa3:86-88, 192).
    The summation order is numpy's pairwise order; the engine's differs (documented
    tolerance 1e-10 relative, observed ~1e-15)."""
    level = h.max_level if level is None else level
    total = 0.0
    for i in range(h.n_fields):
        r = residual_interior(h, level, i)
        total = total + np.sum(r * r)
    return abs(np.sqrt(total))


def mg_solve(h: Hierarchy, ops, settings: SolverSettings, infinity=1e100, solver_iteration_limit=None):
    """Outer iteration of the generated solver; returns
    (residual history [r0..rn], convergence_factor, n_iterations) with the conventions of
    ProgramGenerator.parse_output (exastencils.py:539-584)."""
    history = [residual_norm(h)]
    factors = []
    n = 0
    hit_limit = True
    for n in range(1, settings.max_iterations + 1):
        run_ops(h, ops, settings)
        r = residual_norm(h)
        with np.errstate(all="ignore"):
            factors.append(r / history[-1] if history[-1] != 0 else float("nan"))
        history.append(r)
        if not np.isfinite(r):
            break
        if r < settings.target_reduction * history[0]:
            hit_limit = False
            break
    return history, *summarise(history, factors, n, hit_limit, infinity, solver_iteration_limit)


def summarise(history, factors, n, hit_limit, infinity=1e100, solver_iteration_limit=None):
    rho_inf = math.sqrt(infinity)
    count = 0
    cleaned = []
    for rho in factors:
        if math.isinf(rho) or math.isnan(rho):
            cleaned.append(rho_inf)
        else:
            cleaned.append(rho)
            count += 1
    if count > 0:
        factor = 1.0
        exponent = 1.0 / len(cleaned)
        for rho in cleaned:
            factor *= math.pow(rho, exponent)
    else:
        factor = infinity
    iterations = float(n)
    if hit_limit or n == 0:
        iterations = infinity
    if solver_iteration_limit is not None and iterations >= solver_iteration_limit:
        iterations = infinity
    return factor, iterations


# ----------------------------------------------------------------------------------------
# preconditioned BiCGStab outer solver of the Helmholtz problem
# ----------------------------------------------------------------------------------------

def helmholtz_bicgstab_solve(h: Hierarchy, ops, settings: SolverSettings, operator_a, rhs, infinity=1e100,
                             solver_iteration_limit=None):
    """PreconditionedBiCGStab@finest, statement by statement (Helmholtz/...exa3:144-200).
    `operator_a` = unshifted Helmholtz table on the finest level, the cycle `ops`
    approximately inverts the shifted operator stored in `h.operators`."""
    L = h.max_level
    dt = h.dtype
    shape = (2 ** L + 1,) * h.dimension
    solution = np.zeros(shape, dt)
    h.apply_bc(L, [solution])

    def A(u):
        out = np.zeros(shape, dt)
        _interior(out)[...] = apply_operator(operator_a, [u], 0)
        return out

    def dot(a, b):
        return grid_dot(a, b)

    def precondition(v):
        h.sol[L][0][...] = 0
        h.rhs[L][0][...] = v
        run_ops(h, ops, settings)
        return h.sol[L][0]

    residual = np.zeros(shape, dt)
    _interior(residual)[...] = _interior(rhs) - _interior(A(solution))
    cur = root(dot(residual, residual))
    init = cur
    history = [modulus(cur)]
    if cur == 0.0:
        return history, infinity, infinity
    alpha = beta = rho_new = omega = 1.0 + 0j
    res_hat = residual.copy()
    ap = np.zeros(shape, dt)
    p = np.zeros(shape, dt)
    max_its = settings.max_iterations
    converged_at = None
    for step in range(max_its):
        rho = rho_new
        rho_new = dot(res_hat, residual)
        beta = mul(div(rho_new, rho), div(alpha, omega))
        p = residual + mul(beta, p - mul(omega, ap))
        u = precondition(p)
        ap = A(u)
        alpha = div(rho_new, dot(res_hat, ap))
        hh = solution + mul(alpha, u)
        s = residual - mul(alpha, ap)
        u = precondition(s)
        t = A(u)
        omega = div(dot(t, s), dot(t, t))
        solution = hh + mul(omega, u)
        h.apply_bc(L, [solution])
        residual = s - mul(omega, t)
        cur = root(dot(residual, residual))
        history.append(modulus(cur))
        if not math.isfinite(history[-1]):
            break
        if history[-1] < settings.target_reduction * modulus(init):
            converged_at = step
            break
    # output contract: one "convergence factor" line = |cur|/|init|; iterations = curStep
    # (0-based loop counter, exa3:165, 192-195); the limit message maps to infinity.
    factor = history[-1] / history[0]
    if not math.isfinite(factor):
        factor = math.sqrt(infinity)
    if converged_at is None:
        iterations = infinity
    else:
        iterations = float(converged_at) if converged_at > 0 else infinity
    if solver_iteration_limit is not None and iterations >= solver_iteration_limit:
        iterations = infinity
    return history, factor, iterations
