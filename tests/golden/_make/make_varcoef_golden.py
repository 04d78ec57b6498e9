"""Generate tests/golden/varcoef_golden.json from the UNMODIFIED reference stencil gallery.

Run in the build container only (needs /root/reference):

    python tests/golden/_make/make_varcoef_golden.py

`Poisson2DVariableCoefficients(get_coefficient_2D, position).generate_stencil(grid)` and the 3D class
(evostencils/stencils/gallery.py:87-185) give the variable-coefficient stencil at ONE node position;
the generated solver evaluates the same expressions (`generate_exa3`, :113-131, :163-185) at every node.
The fixture holds the reference's values at a spread of node positions and levels (boundary-adjacent
nodes, the centre, asymmetric positions); tests/test_variable_coefficients.py compares the oracle's
per-node tables and the host-side stencil generator against them."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.abspath(os.path.join(HERE, "..", "..", ".."))
sys.path.insert(0, "/root/reference")

from evostencils.stencils import gallery  # noqa: E402


class _Grid:
    def __init__(self, spacing):
        self.spacing = spacing


def main():
    cases = []
    for dim, levels in ((2, (2, 3, 5, 8, 12)), (3, (2, 3, 5, 9))):
        for level in levels:
            n = 2 ** level + 1
            h = 1.0 / 2 ** level
            picks = sorted({1, 2, n // 4, n // 2, (3 * n) // 4 + 1, n - 3, n - 2} & set(range(1, n - 1)))
            nodes = []
            for k, i in enumerate(picks):
                node = [i, picks[(k * 2 + 1) % len(picks)]] + ([picks[(k * 3 + 2) % len(picks)]] if dim == 3 else [])
                nodes.append(node)
            for node in nodes:
                position = [i * h for i in node]
                if dim == 2:
                    gen = gallery.Poisson2DVariableCoefficients(gallery.get_coefficient_2D, position)
                else:
                    gen = gallery.Poisson3DVariableCoefficients(gallery.get_coefficient_3D, position)
                stencil = gen.generate_stencil(_Grid((h,) * dim))
                cases.append({"dimension": dim, "level": level, "node": node,
                              "entries": [[list(off), float(val).hex()] for off, val in stencil.entries]})
    out = os.path.join(REPO, "tests", "golden", "varcoef_golden.json")
    with open(out, "w") as f:
        json.dump({"source": "evostencils/stencils/gallery.py:87-185 (kappa = 10)", "cases": cases}, f, indent=0)
    print(len(cases), "cases ->", out)


if __name__ == "__main__":
    main()
