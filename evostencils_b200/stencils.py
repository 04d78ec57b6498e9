"""Stencil algebra and stencil gallery for the B200 multigrid engine.

Host-side mirror of the reference's stencil layer:
  * constant stencils  -> /root/reference/evostencils/stencils/constant.py:1-149
  * periodic stencils  -> /root/reference/evostencils/stencils/multiple.py:4-240
  * gallery            -> /root/reference/evostencils/stencils/gallery.py:16-219

Design differences (this is not a translation):
  * a `Stencil` is an ordered offset->value map (dict) instead of a tuple list that
    is scanned linearly on every combine;
  * a `PeriodicStencil` stores one `Stencil` per residue class in a flat dict keyed
    by the index tuple, with an explicit `period`;
  * the full-weighting / multilinear inter-grid weights are closed-form tensor
    products here (the reference obtains them from `lfa_lab`, which is a
    third-party C++/SWIG package, gallery.py:193-198, 210-215).

Both classes keep the *duck-typed* attributes the reference's consumers read
(`entries`, `dimension`, `constant_stencils`) so that `lowering.py` can take either a
reference IR graph or one built from `evostencils_b200.ir`.

Offsets follow the reference convention: component 0 is the x direction (`i0`).
"""
from __future__ import annotations

import itertools
import math
from typing import Callable, Dict, Iterable, Sequence, Tuple

Offset = Tuple[int, ...]


class Stencil:
    """Constant-coefficient stencil: ordered list of (offset, value)."""

    __slots__ = ("_map", "_dimension")

    def __init__(self, entries: Iterable[Tuple[Offset, object]] = (), dimension: int | None = None):
        self._map: Dict[Offset, object] = {}
        for off, val in entries:
            off = tuple(int(o) for o in off)
            if off in self._map:
                self._map[off] = self._map[off] + val
            else:
                self._map[off] = val
        if dimension is None:
            if not self._map:
                raise ValueError("dimension required for an empty stencil")
            dimension = len(next(iter(self._map)))
        self._dimension = dimension

    # -- duck-typed surface shared with the reference ---------------------------------
    @property
    def entries(self) -> Tuple[Tuple[Offset, object], ...]:
        return tuple(self._map.items())

    @property
    def dimension(self) -> int:
        return self._dimension

    @property
    def number_of_entries(self) -> int:
        return len(self._map)

    # -- algebra -----------------------------------------------------------------------
    def get(self, offset: Offset, default=0.0):
        return self._map.get(tuple(offset), default)

    def map(self, f: Callable[[Offset, object], Tuple[Offset, object]]) -> "Stencil":
        return Stencil((f(o, v) for o, v in self._map.items()), self._dimension)

    def filter(self, pred: Callable[[Offset, object], bool]) -> "Stencil":
        return Stencil(((o, v) for o, v in self._map.items() if pred(o, v)), self._dimension)

    def diagonal(self) -> "Stencil":
        return self.filter(lambda o, v: not any(o))

    def scaled(self, factor) -> "Stencil":
        return self.map(lambda o, v: (o, factor * v))

    def transposed(self) -> "Stencil":
        return self.map(lambda o, v: (tuple(-c for c in o), v))

    def combined(self, other: "Stencil", f) -> "Stencil":
        # entries of `self` keep their position; new offsets of `other` are appended
        # (same ordering rule as constant.py:54-67)
        out = dict(self._map)
        for o, v in other._map.items():
            out[o] = f(out[o], v) if o in out else v
        return Stencil(out.items(), self._dimension)

    def __add__(self, other):
        return self.combined(other, lambda a, b: a + b)

    def __sub__(self, other):
        # constant.py:117-118 -- an offset only present in `other` is appended unnegated
        # there; here it is negated, which is the mathematically correct difference.
        out = dict(self._map)
        for o, v in other._map.items():
            out[o] = out[o] - v if o in out else -v
        return Stencil(out.items(), self._dimension)

    def __matmul__(self, other: "Stencil") -> "Stencil":
        """Operator product self*other (convolution of offsets, constant.py:125-142)."""
        out: Dict[Offset, object] = {}
        for o2, v2 in other._map.items():
            for o1, v1 in self._map.items():
                o = tuple(a + b for a, b in zip(o1, o2))
                out[o] = out[o] + v1 * v2 if o in out else v1 * v2
        return Stencil(out.items(), self._dimension)

    def __repr__(self):
        return f"Stencil({self.entries!r})"


def unit_stencil(dimension: int) -> Stencil:
    return Stencil([((0,) * dimension, 1.0)], dimension)


def null_stencil(dimension: int) -> Stencil:
    return Stencil((), dimension)


class PeriodicStencil:
    """One constant stencil per residue class `index % period` (multiple.py:4-17)."""

    def __init__(self, period: Sequence[int], table: Dict[Offset, Stencil], dimension: int):
        self.period = tuple(int(p) for p in period)
        self.table = dict(table)
        self._dimension = dimension
        assert len(self.period) == dimension
        for idx in itertools.product(*(range(p) for p in self.period)):
            assert idx in self.table, f"missing residue class {idx}"

    @property
    def dimension(self):
        return self._dimension

    @property
    def constant_stencils(self):
        """Nested tuples, nesting level k <-> offset component k (multiple.py:90-101)."""

        def rec(prefix, k):
            if k == self._dimension:
                return self.table[prefix]
            return tuple(rec(prefix + (i,), k + 1) for i in range(self.period[k]))

        return rec((), 0)


def block_diagonal(stencil: Stencil, block_size: Sequence[int]) -> PeriodicStencil:
    """Non-overlapping tiles anchored at index 0: the class `idx` keeps the offsets that
    stay inside the tile, 0 <= idx+o < block_size (multiple.py:204-217)."""
    dim = stencil.dimension
    assert len(block_size) == dim
    table = {}
    for idx in itertools.product(*(range(b) for b in block_size)):
        table[idx] = stencil.filter(
            lambda o, v, idx=idx: all(0 <= i + c < b for i, c, b in zip(idx, o, block_size)))
    return PeriodicStencil(block_size, table, dim)


def periodic_shape(stencil) -> Tuple[int, ...]:
    """Period of a (possibly reference-typed) periodic stencil; () for a constant one."""
    if hasattr(stencil, "period"):
        return tuple(stencil.period)
    if not hasattr(stencil, "constant_stencils"):
        return ()
    shape = []
    level = stencil.constant_stencils
    for _ in range(stencil.dimension):
        shape.append(len(level))
        level = level[0]
    return tuple(shape)


# --------------------------------------------------------------------------------------
# gallery
# --------------------------------------------------------------------------------------

def poisson(spacing: Sequence[float]) -> Stencil:
    """(2d+1)-point negative Laplacian, values as in gallery.py:32-44 / 60-71 and the
    `.exa2` operator blocks (example_problems/Poisson/2D_FD_Poisson_fromL2.exa2:9-15)."""
    dim = len(spacing)
    centre = 0.0
    entries = []
    for d, h in enumerate(spacing):
        centre = centre + 2 / (h * h)
    entries.append(((0,) * dim, centre))
    for d, h in enumerate(spacing):
        for s in (-1, 1):
            off = [0] * dim
            off[d] = s
            entries.append((tuple(off), -1 / (h * h)))
    return Stencil(entries, dim)


def second_derivative(spacing: Sequence[float], axis: int) -> Stencil:
    """`dxx`/`dyy` of LinearElasticity/2D_FD_LinearElasticity_fromL2.exa2:11-21."""
    dim = len(spacing)
    h = spacing[axis]
    lo = [0] * dim
    hi = [0] * dim
    lo[axis], hi[axis] = -1, 1
    return Stencil([((0,) * dim, -2.0 / (h ** 2)), (tuple(lo), 1.0 / (h ** 2)), (tuple(hi), 1.0 / (h ** 2))], dim)


def laplace(spacing: Sequence[float]) -> Stencil:
    """Positive Laplacian of the elasticity example (.exa2:23-29)."""
    dim = len(spacing)
    centre = 0.0
    first = True
    for h in spacing:
        centre = (-2.0 / (h ** 2)) if first else centre - 2.0 / (h ** 2)
        first = False
    entries = [((0,) * dim, centre)]
    for d, h in enumerate(spacing):
        for s in (-1, 1):
            off = [0] * dim
            off[d] = s
            entries.append((tuple(off), 1.0 / (h ** 2)))
    return Stencil(entries, dim)


def mixed_derivative_2d(spacing: Sequence[float]) -> Stencil:
    """`dxy` 4-point corner stencil (.exa2:31-36)."""
    hx, hy = spacing
    w = 1.0 / (4 * hx * hy)
    return Stencil([((-1, 1), -w), ((1, 1), w), ((-1, -1), w), ((1, -1), -w)], 2)


def helmholtz_2d(spacing: Sequence[float], k: float, shift: complex = 1.0) -> Stencil:
    """`A` (shift=1) and `M` (shift=1+0.5j) of Helmholtz/2D_FD_Helmholtz_fromL3.exa3:55-69."""
    hx, hy = spacing
    centre = 2.0 / (hx ** 2) + 2.0 / (hy ** 2) - k ** 2 * shift
    return Stencil([((0, 0), centre), ((-1, 0), -1.0 / (hx ** 2)), ((1, 0), -1.0 / (hx ** 2)),
                    ((0, -1), -1.0 / (hy ** 2)), ((0, 1), -1.0 / (hy ** 2))], 2)


def variable_coefficient_kappa(*pos: float) -> float:
    """exp(kappa * prod(x_d - x_d^2)), kappa = 10 (gallery.py:87-90, 124-127)."""
    p = 1.0
    first = True
    for x in pos:
        t = x - (x * x)
        p = t if first else p * t
        first = False
    return math.exp(10.0 * p)


def variable_coefficient_exp(kappa: float, *pos: float) -> float:
    """get_coefficient_2D / get_coefficient_3D (gallery.py:87-90, 137-140) for any kappa:
    exp(kappa * ((x - x*x) * (y - y*y) [* (z - z*z)])), products grouped left to right."""
    p = None
    for x in pos:
        t = x - (x * x)
        p = t if p is None else p * t
    return math.exp(kappa * p)


def variable_poisson(spacing: Sequence[float], position: Sequence[float], kappa: float = 10.0) -> Stencil:
    """Stencil of -div(c grad u) at `position`: Poisson2DVariableCoefficients.generate_stencil /
    Poisson3DVariableCoefficients.generate_stencil (gallery.py:93-111, 143-161): c at the half-grid
    positions, off-centre value (-1.0 * c) / (h*h), centre = sum over the axes of (c+ + c-) / (h*h),
    accumulated axis by axis.  The reference evaluates this at ONE position (for its convergence
    model); the generated solver evaluates it at every node (`Operator A from Stencil`, :113-131), which
    is what the engine does (csrc/mg_kernels_generic.cuh vc_apply)."""
    d = len(spacing)
    assert len(position) == d

    def face(axis, sign):
        q = list(position)
        q[axis] = q[axis] + sign * (0.5 * spacing[axis])
        return variable_coefficient_exp(kappa, *q)

    centre = None
    entries = []
    for axis in range(d):
        hh = spacing[axis] * spacing[axis]
        cp, cm = face(axis, +1), face(axis, -1)
        term = (cp + cm) / hh
        centre = term if centre is None else centre + term
        for sign, c in ((+1, cp), (-1, cm)):
            off = [0] * d
            off[axis] = sign
            entries.append((tuple(off), (-1.0 * c) / hh))
    entries.append(((0,) * d, centre))
    return Stencil(entries, d)


def full_weighting(dimension: int) -> Stencil:
    """Full-weighting restriction, tensor product of [1/4, 1/2, 1/4]; the coarse node I
    gathers from fine nodes 2I+o (reference DSL form `[i] from [2*i + o]`,
    exastencils.py:1126-1147; generator class gallery.py:205-219)."""
    w1 = {-1: 0.25, 0: 0.5, 1: 0.25}
    entries = []
    for o in itertools.product((-1, 0, 1), repeat=dimension):
        v = 1.0
        for c in o:
            v *= w1[c]
        entries.append((o, v))
    return Stencil(entries, dimension)


def multilinear_interpolation(dimension: int) -> Stencil:
    """Multilinear prolongation, tensor product of [1/2, 1, 1/2] (gallery.py:188-202;
    DSL form `[i] from [0.5*(i + o)]`, exastencils.py:1103-1124)."""
    w1 = {-1: 0.5, 0: 1.0, 1: 0.5}
    entries = []
    for o in itertools.product((-1, 0, 1), repeat=dimension):
        v = 1.0
        for c in o:
            v *= w1[c]
        entries.append((o, v))
    return Stencil(entries, dimension)


def canonical_entries(stencil) -> Tuple[Tuple[Offset, object], ...]:
    """Entries sorted in memory order (slowest axis first, x fastest), zeros dropped.

    This is the accumulation order shared by the oracle and the CUDA kernels
    (decision O3'); it makes the arithmetic independent of how the stencil
    was assembled."""
    items = [(tuple(int(c) for c in o), v) for o, v in stencil.entries if v != 0]
    items.sort(key=lambda e: tuple(reversed(e[0])))
    return tuple(items)
